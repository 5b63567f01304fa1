"""NVTX ranges for the profilers' timelines (the reference has only ``time.time()`` brackets).

``with nvtx_range("epoch 3"):`` shows up as a named range in Nsight Systems / Nsight Compute
(``ncu --nvtx --nvtx-include "train/"`` restricts a capture to the training part of an epoch). Without a
CUDA device, or if the NVTX library cannot be loaded, the ranges are no-ops.
"""
from __future__ import annotations

import contextlib

_push = _pop = None
_state = {"ready": False}


def _init() -> None:
    global _push, _pop
    _state["ready"] = True
    try:
        import torch

        if not torch.cuda.is_available():
            return
        from torch.cuda import nvtx

        nvtx.range_push("dnnb200/nvtx-probe")
        nvtx.range_pop()
        _push, _pop = nvtx.range_push, nvtx.range_pop
    except Exception:       # library missing: stay silent, ranges are an aid, not a dependency
        _push = _pop = None


def enabled() -> bool:
    if not _state["ready"]:
        _init()
    return _push is not None


@contextlib.contextmanager
def nvtx_range(name: str):
    if not _state["ready"]:
        _init()
    if _push is None:
        yield
        return
    _push(name)
    try:
        yield
    finally:
        _pop()
