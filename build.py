#!/usr/bin/env python
"""Build the native runtime IN-TREE for sm_100a with nvcc (no torch headers needed).

    python build.py            # -> distributed_neural_network_b200/_native/libdnnb200.so
    python build.py --force    # rebuild everything
    python build.py --sass     # also dump SASS listings into profiles/sass/

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import argparse
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
OUT_DIR = ROOT / "distributed_neural_network_b200" / "_native"
# experiments: DNN_BUILD_TAG=x DNN_BUILD_DEFINES="-DFOO=1 -DBAR=2" python build.py -> libdnnb200_x.so (load it with
# DNNB200_LIB=<path>); the default build has neither
_TAG = os.environ.get("DNN_BUILD_TAG", "")
OBJ_DIR = OUT_DIR / ("obj" + ("_" + _TAG if _TAG else ""))
LIB = OUT_DIR / ("libdnnb200" + ("_" + _TAG if _TAG else "") + ".so")

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--use_fast_math", "-Xptxas", "-v",
          "-I", str(ROOT / "include"), "-I", str(CSRC)] + os.environ.get("DNN_BUILD_DEFINES", "").split()


def sources():
    return sorted(CSRC.glob("*.cu"))


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(ARCH + COMMON).encode())
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()


def compile_one(src: Path, force: bool) -> Path:
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha")
    dig = _digest(src)
    if not force and obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj
    cmd = [NVCC, *ARCH, *COMMON, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    (OBJ_DIR / (src.stem + ".ptxas.log")).write_text(r.stderr)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        if LIB.exists():
            LIB.unlink()  # never leave a stale library behind: a GPU run must not silently test old code
        raise RuntimeError(f"nvcc failed on {src.name}")
    stamp.write_text(dig)
    return obj


def build(force: bool = False, verbose: bool = True) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: compile_one(s, force), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if force or not LIB.exists() or LIB.stat().st_mtime < newest:
        cmd = [NVCC, *ARCH, "-shared", "-o", str(LIB), *map(str, objs)]  # static cudart (nvcc default); driver API is resolved at run time
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    if verbose:
        print(f"built {LIB} ({LIB.stat().st_size >> 10} KiB) from {len(srcs)} sources")
    return LIB


def dump_sass() -> None:
    out = ROOT / "profiles" / "sass"
    out.mkdir(parents=True, exist_ok=True)
    for src in sources():
        obj = OBJ_DIR / (src.stem + ".o")
        r = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True)
        (out / (src.stem + ".sass")).write_text(r.stdout)
        log = OBJ_DIR / (src.stem + ".ptxas.log")
        if log.exists():
            (out / (src.stem + ".ptxas.txt")).write_text(log.read_text())


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--sass", action="store_true")
    a = ap.parse_args()
    build(force=a.force)
    if a.sass:
        dump_sass()
