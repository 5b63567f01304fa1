"""Command lines of the three entry points.

Same flags, spellings, dests and defaults as the reference's argparse blocks
(data_parallelism_train.py:259-272, model_replication_train.py:153-160), with the missing
``type=`` added -- in the reference, values given on the command line reach SGD and
DataLoader as strings (quirk Q5). Everything below the separator is new.
"""
from __future__ import annotations

import argparse

from ..trainer import TrainConfig


def add_reference_flags(p: argparse.ArgumentParser, epochs_default: int, with_dp_flags: bool) -> None:
    p.add_argument("--lr", dest="lr", default=0.001, type=float)
    p.add_argument("--momentum", dest="momentum", default=0.9, type=float)
    p.add_argument("--batch-size", dest="bs", default=16, type=int)
    p.add_argument("--epochs", dest="epochs", default=epochs_default, type=int)
    if with_dp_flags:
        p.add_argument("--nb-proc", dest="nb_proc", default=4, type=int)
        p.add_argument("--failure-probability", dest="failure_probability", default=0.0, type=float,
                       help="Probability of simulated process failure at each epoch")
        p.add_argument("--failure-duration", dest="failure_duration", default=0.0, type=float,
                       help="Duration of simulated process failure in seconds")


def add_new_flags(p: argparse.ArgumentParser, sync_default: str) -> None:
    g = p.add_argument_group("B200-native additions")
    g.add_argument("--sync-mode", default=sync_default,
                   choices=["step-allreduce", "epoch-average", "replication", "none"],
                   help="step-allreduce: fused per-step gradient all-reduce (default for data parallelism); "
                        "epoch-average: the reference's once-per-epoch weight averaging")
    g.add_argument("--reference-topology", action="store_true",
                   help="rank 0 is a parent that only evaluates; ranks 1..P-1 train (reference roles)")
    g.add_argument("--failure-mode", default="sleep", choices=["sleep", "drop"],
                   help="sleep: the failing rank sleeps and everyone waits (reference); "
                        "drop: it leaves the peer group and the survivors keep training")
    g.add_argument("--device", default="auto", choices=["auto", "cpu", "cuda"])
    g.add_argument("--data-root", default="./data")
    g.add_argument("--synthetic", action="store_true", help="CIFAR-10-shaped synthetic data (no files needed)")
    g.add_argument("--train-size", type=int, default=None)
    g.add_argument("--test-size", type=int, default=None)
    g.add_argument("--seed", type=int, default=None)
    g.add_argument("--log-dir", default="log")
    g.add_argument("--jsonl", default=None, help="per-epoch metrics, one JSON object per line")
    g.add_argument("--save", default=None, help="write the state_dict (reference format) here after every epoch")
    g.add_argument("--resume", default=None, help="load a state_dict (plus the .resume sidecar if present)")
    g.add_argument("--input-pipeline", default="device", choices=["device", "host"])
    g.add_argument("--allreduce-variant", default="auto", choices=["auto", "one-shot", "one-shot-128", "two-shot", "nvls"],
                   help="gradient exchange: one-shot push, two-shot (owner-reduce + broadcast), chosen by world size, or "
                        "(parameter averaging only) nvls = in-switch reduction over NVLink multicast")
    g.add_argument("--step-path", default="fused", choices=["fused", "autograd"],
                   help="fused: the two-kernel step engine; autograd: layer kernels through torch autograd with the "
                        "gradient exchanged per bucket on a side stream while backward runs (large batches)")
    g.add_argument("--precision", default=None, choices=["fp32", "tf32", "tf32-dgrad"],
                   help="convolution stages of the fused GPU step: fp32 = FFMA throughout (default: parity with PyTorch "
                        "fp32, and measured faster at this model's GEMM sizes), tf32 = tcgen05 tensor cores")
    g.add_argument("--neptune", action="store_true", help="log to Neptune (needs NEPTUNE_API_TOKEN in the environment)")
    g.add_argument("--quiet", action="store_true")


def config_from_args(a: argparse.Namespace) -> TrainConfig:
    return TrainConfig(
        lr=a.lr, momentum=a.momentum, batch_size=a.bs, epochs=a.epochs,
        nb_proc=getattr(a, "nb_proc", 1),
        failure_probability=getattr(a, "failure_probability", 0.0),
        failure_duration=getattr(a, "failure_duration", 0.0),
        failure_mode=a.failure_mode, sync_mode=a.sync_mode, reference_topology=a.reference_topology,
        device=a.device, data_root=a.data_root, synthetic=True if a.synthetic else None,
        train_size=a.train_size, test_size=a.test_size, seed=a.seed, log_dir=a.log_dir, jsonl=a.jsonl,
        save=a.save, resume=a.resume, input_pipeline=a.input_pipeline, allreduce_variant=a.allreduce_variant,
        neptune=a.neptune, quiet=a.quiet, precision=a.precision, step_path=a.step_path)
