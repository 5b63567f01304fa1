#!/usr/bin/env python
"""Data-parallel CIFAR-10 CNN training -- the core entry point.

    mpiexec -n 4 python data_parallelism_train.py --nb-proc 4 --batch-size 16      # with an MPI runtime
    bin/mpiexec -n 4 python data_parallelism_train.py --nb-proc 4 --batch-size 16  # bundled spawner
    torchrun --nproc-per-node 4 data_parallelism_train.py --nb-proc 4

Same name, flags and log files as the reference script (data_parallelism_train.py:56-272). The
default synchronisation is the fused per-step gradient all-reduce over NVLink peer memory;
``--sync-mode epoch-average --reference-topology`` reproduces the reference's algorithm and roles
(rank 0 parent, once-per-epoch weight averaging).
"""
import argparse
import sys

from distributed_neural_network_b200.trainer import Trainer
from distributed_neural_network_b200.utils import cli


def main(args) -> int:
    cfg = cli.config_from_args(args)
    trainer = Trainer(cfg)
    try:
        trainer.fit()
    finally:
        trainer.close()
        trainer.comm.close()
    return 0


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    cli.add_reference_flags(parser, epochs_default=25, with_dp_flags=True)
    cli.add_new_flags(parser, sync_default="step-allreduce")
    sys.exit(main(parser.parse_args()))
