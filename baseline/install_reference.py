"""Put the UNMODIFIED reference under baseline/_ref (git-ignored; it travels to the GPU box with the snapshot).

`pip install --target baseline/_ref /root/reference` cannot work: the reference is four loose scripts with neither
setup.py nor pyproject.toml. So its files are copied verbatim. Called by
`__graft_entry__.build()` and by `bench.py --impl reference`, so a fresh checkout of this repository on a machine
that has the read-only reference at /root/reference gets the copy back without a manual step.
"""
from __future__ import annotations

import shutil
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = Path("/root/reference")
DST = HERE / "_ref"
FILES = ["data_parallelism_train.py", "model_replication_train.py", "single_proc_train.py", "run_training.sh",
         "README.md", "LICENSE"]


def ensure() -> bool:
    """True if baseline/_ref holds the reference's scripts afterwards."""
    if (DST / "data_parallelism_train.py").exists() and (DST / "models" / "model.py").exists():
        return True
    if not (SRC / "data_parallelism_train.py").exists():
        return False
    DST.mkdir(parents=True, exist_ok=True)
    for f in FILES:
        if (SRC / f).exists():
            shutil.copy2(SRC / f, DST / f)
    if (DST / "models").exists():
        shutil.rmtree(DST / "models")
    shutil.copytree(SRC / "models", DST / "models", ignore=shutil.ignore_patterns("__pycache__"))
    return True


if __name__ == "__main__":
    print("baseline/_ref:", "ready" if ensure() else "reference checkout not found at /root/reference")
