"""Flat parameter layout shared by Python and the CUDA kernels.

The ten tensors of the CNN (reference: models/model.py:13-18) live in ONE fp32
buffer so that the gradient exchange, the optimizer and the checkpoint code all
see a single contiguous vector. Every tensor starts on a 128-byte boundary
(32 floats) so kernels can use 16-byte vector loads; the gaps stay zero forever
(their gradient and momentum are never written).

The same table is compiled into csrc/layout.h -- tests/test_flat_and_model.py checks that
the two agree.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Tuple

import torch

ALIGN = 32  # floats (128 B)

# name -> shape, in state_dict order (reference: models/model.py:13-18)
PARAM_SHAPES: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict(
    [
        ("conv1.weight", (6, 3, 5, 5)),
        ("conv1.bias", (6,)),
        ("conv2.weight", (16, 6, 5, 5)),
        ("conv2.bias", (16,)),
        ("fc1.weight", (120, 400)),
        ("fc1.bias", (120,)),
        ("fc2.weight", (84, 120)),
        ("fc2.bias", (84,)),
        ("fc3.weight", (10, 84)),
        ("fc3.bias", (10,)),
    ]
)


def _numel(shape: Tuple[int, ...]) -> int:
    n = 1
    for s in shape:
        n *= s
    return n


def _build() -> Tuple[Dict[str, int], int]:
    offs: Dict[str, int] = {}
    cur = 0
    for name, shape in PARAM_SHAPES.items():
        offs[name] = cur
        cur += _numel(shape)
        cur = (cur + ALIGN - 1) // ALIGN * ALIGN
    return offs, cur


PARAM_OFFSETS, FLAT_NUMEL = _build()
PARAM_NUMEL = sum(_numel(s) for s in PARAM_SHAPES.values())  # 62006 real parameters

assert PARAM_NUMEL == 62006
assert FLAT_NUMEL == 62144


def views(flat: torch.Tensor) -> "OrderedDict[str, torch.Tensor]":
    """Per-tensor views (no copy) into a flat buffer of FLAT_NUMEL floats."""
    assert flat.numel() == FLAT_NUMEL and flat.dim() == 1
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for name, shape in PARAM_SHAPES.items():
        off = PARAM_OFFSETS[name]
        out[name] = flat[off : off + _numel(shape)].view(shape)
    return out


def pack(state_dict, out: torch.Tensor | None = None) -> torch.Tensor:
    """state_dict (10 keys) -> flat fp32 buffer."""
    if out is None:
        any_t = next(iter(state_dict.values()))
        out = torch.zeros(FLAT_NUMEL, dtype=torch.float32, device=any_t.device)
    v = views(out)
    for name in PARAM_SHAPES:
        v[name].copy_(state_dict[name])
    return out


def unpack(flat: torch.Tensor, clone: bool = True) -> "OrderedDict[str, torch.Tensor]":
    """flat buffer -> OrderedDict with the reference's 10 keys/shapes."""
    v = views(flat)
    if clone:
        return OrderedDict((k, t.detach().clone()) for k, t in v.items())
    return v


def real_mask(device=None) -> torch.Tensor:
    """Boolean mask of the positions that hold real parameters (not padding)."""
    m = torch.zeros(FLAT_NUMEL, dtype=torch.bool, device=device)
    for name, shape in PARAM_SHAPES.items():
        off = PARAM_OFFSETS[name]
        m[off : off + _numel(shape)] = True
    return m
