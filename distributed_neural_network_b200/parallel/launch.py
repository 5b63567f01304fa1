"""mpiexec-compatible process spawner for boxes without an MPI runtime.

    python -m distributed_neural_network_b200.parallel.launch -n 4 python data_parallelism_train.py --nb-proc 4
    bin/mpiexec -n 4 python data_parallelism_train.py --nb-proc 4        # same thing

keeps the reference's documented command line (reference: README.md:27-29, run_training.sh)
working where ``mpiexec``/``mpi4py`` are not installed. One process per rank (= per GPU when
CUDA is present), RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT in the environment,
rendezvous on 127.0.0.1. If one rank dies the others are terminated and its exit code is returned.
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def spawn(n: int, cmd: List[str], env_extra=None, gpus: int | None = None) -> int:
    port = _free_port()
    procs: List[subprocess.Popen] = []
    if gpus is None:
        try:
            import torch

            gpus = torch.cuda.device_count()
        except Exception:
            gpus = 0
    for rank in range(n):
        env = dict(os.environ)
        env.update({
            "RANK": str(rank), "WORLD_SIZE": str(n), "LOCAL_RANK": str(rank), "LOCAL_WORLD_SIZE": str(n),
            "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
            "DNNB200_LAUNCHER": "1",
        })
        if gpus:
            env["DNNB200_DEVICE"] = str(rank % gpus)
        if env_extra:
            env.update(env_extra)
        procs.append(subprocess.Popen(cmd, env=env, start_new_session=True))
    rc = 0
    try:
        alive = set(range(n))
        while alive:
            for r in list(alive):
                code = procs[r].poll()
                if code is None:
                    continue
                alive.discard(r)
                if code != 0 and rc == 0:
                    rc = code
                    for o in alive:  # one rank failed: stop the rest (exact pids we started)
                        try:
                            os.killpg(procs[o].pid, signal.SIGTERM)
                        except ProcessLookupError:
                            pass
            time.sleep(0.02)
    except KeyboardInterrupt:
        for p in procs:
            try:
                os.killpg(p.pid, signal.SIGTERM)
            except ProcessLookupError:
                pass
        rc = 130
    return rc


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="mpiexec", description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("-n", "-np", "--np", "-c", dest="n", type=int, required=True, help="number of processes")
    # what people have in their Open MPI / MPICH command lines; accepted so that those lines keep working
    ap.add_argument("--oversubscribe", "--allow-run-as-root", "--use-hwthread-cpus", "--report-bindings",
                    "--tag-output", dest="_ignored", action="store_true", help="accepted, no effect (one node, one spawner)")
    ap.add_argument("--bind-to", "--map-by", "--rank-by", dest="_ignored_1", metavar="X",
                    help="accepted with its value, no effect")
    ap.add_argument("--mca", dest="_ignored_2", nargs=2, action="append", metavar=("KEY", "VALUE"),
                    help="accepted with its two values, no effect")
    ap.add_argument("-x", dest="export", action="append", default=[], metavar="VAR[=VALUE]",
                    help="Open MPI: export VAR (with VALUE, or its current value) to the processes")
    ap.add_argument("-genv", "-env", dest="genv", action="append", nargs=2, default=[], metavar=("VAR", "VALUE"),
                    help="MPICH: set VAR=VALUE for the processes")
    ap.add_argument("cmd", nargs=argparse.REMAINDER, help="program and its arguments")
    a = ap.parse_args(argv)
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    if not cmd:
        ap.error("no command given")
    env_extra = {}
    for item in a.export:
        name, eq, value = item.partition("=")
        env_extra[name] = value if eq else os.environ.get(name, "")
    for name, value in a.genv:
        env_extra[name] = value
    return spawn(a.n, cmd, env_extra=env_extra or None)


if __name__ == "__main__":
    sys.exit(main())
