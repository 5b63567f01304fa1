"""Checkpoint / resume in the reference's state_dict format.

The reference never writes a checkpoint (no torch.save anywhere; the state_dict is only
its wire format, data_parallelism_train.py:118,210 -- SURVEY fact 4). Here ``--save`` writes
exactly that object -- an OrderedDict with the ten keys/shapes/dtypes of models/model.py:13-18
-- so the file loads into the reference ``Network.load_state_dict`` unchanged. Everything
needed for an EXACT resume (momentum buffer, epoch, step, membership) goes into a sidecar
``<path>.resume`` so the main file stays format-pure.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Any, Dict, Optional

import torch

from ..ops import flat as L


def save(path: str, model, momentum: Optional[torch.Tensor] = None, **meta: Any) -> None:
    sd = OrderedDict((k, v.detach().cpu().clone()) for k, v in model.state_dict().items())
    assert list(sd.keys()) == list(L.PARAM_SHAPES.keys())
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp"
    torch.save(sd, tmp)
    os.replace(tmp, path)  # never leave a torn file behind
    side: Dict[str, Any] = dict(meta)
    if momentum is not None:
        side["momentum"] = momentum.detach().cpu().clone()
    torch.save(side, path + ".resume.tmp")
    os.replace(path + ".resume.tmp", path + ".resume")


def load(path: str, model, momentum: Optional[torch.Tensor] = None) -> Dict[str, Any]:
    """Load weights (and, if the sidecar exists, momentum + metadata). Returns the metadata."""
    sd = torch.load(path, map_location="cpu", weights_only=True)
    model.load_state_dict(sd)
    meta: Dict[str, Any] = {}
    side = path + ".resume"
    if os.path.exists(side):
        meta = torch.load(side, map_location="cpu", weights_only=True)
        m = meta.pop("momentum", None)
        if m is not None and momentum is not None:
            momentum.copy_(m.to(momentum.device))
    return meta
