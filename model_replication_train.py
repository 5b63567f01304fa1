#!/usr/bin/env python
"""Model replication: every trainer sees the FULL training set, weights are averaged per epoch.

    bin/mpiexec -n 4 python model_replication_train.py --batch-size 16 --epochs 10

Same name and flags as the reference script (model_replication_train.py:31-160; --epochs
defaults to 10 there). Rank 0 is the evaluating parent exactly as in the reference unless
``--all-train`` is given.
"""
import argparse
import sys

from distributed_neural_network_b200.trainer import Trainer
from distributed_neural_network_b200.utils import cli


def main(args) -> int:
    cfg = cli.config_from_args(args)
    cfg.sync_mode = "replication"
    cfg.reference_topology = not args.all_train
    trainer = Trainer(cfg)
    try:
        trainer.fit()
    finally:
        trainer.close()
        trainer.comm.close()
    return 0


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    cli.add_reference_flags(parser, epochs_default=10, with_dp_flags=False)
    cli.add_new_flags(parser, sync_default="replication")
    parser.add_argument("--all-train", action="store_true", help="rank 0 trains too (no dedicated parent)")
    sys.exit(main(parser.parse_args()))
