#!/usr/bin/env python
"""Single-process baseline (reference: single_proc_train.py:16-107).

    python single_proc_train.py

The reference hard-codes batch size 4, 15 epochs, lr 0.001, momentum 0.9 and takes no flags;
those are the defaults here, and every one of them can be overridden. Uses ``cuda:0`` when
available (reference :28), through the fused step engine.
"""
import argparse
import sys

from distributed_neural_network_b200.parallel import bootstrap
from distributed_neural_network_b200.trainer import Trainer
from distributed_neural_network_b200.utils import cli


def main(args) -> int:
    cfg = cli.config_from_args(args)
    cfg.sync_mode = "none"
    trainer = Trainer(cfg, comm=bootstrap.Comm())
    print(trainer.device)
    try:
        shard, test = trainer.load_data()
        print(len(shard))
        print(len(test))
        hist = trainer.fit(shard, test)
        for h in hist:
            print(f"Epoch {h['epoch'] + 1}, Average Training Loss: {h['train_loss']:.3f}")
    finally:
        trainer.close()
    return 0


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    parser.add_argument("--lr", dest="lr", default=0.001, type=float)
    parser.add_argument("--momentum", dest="momentum", default=0.9, type=float)
    parser.add_argument("--batch-size", dest="bs", default=4, type=int)
    parser.add_argument("--epochs", dest="epochs", default=15, type=int)
    cli.add_new_flags(parser, sync_default="none")
    sys.exit(main(parser.parse_args()))
