"""The CIFAR-10 CNN with the reference's public API, backed by one flat buffer.

Same constructor, ``forward`` and ``state_dict`` keys/shapes/init as the
reference ``Network`` (reference: models/model.py:9-27), so a checkpoint written
by either loads into the other. What differs is the storage: the ten parameter
tensors are views into ONE fp32 vector (``ops/flat.py``), and so are their
gradients. That is what lets the fused step engine, the peer-memory all-reduce
and the SGD update work on a single 248 KB buffer with no gather/scatter.

On a CUDA device ``forward`` dispatches to the hand-written sm_100a kernels via
``ops/functional.py``; on CPU it runs the plain PyTorch ops (the oracle).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from ..ops import flat as flat_layout
from ..ops import functional as F_native


class _Params(nn.Module):
    """Holds ``weight``/``bias`` Parameters that alias slices of the flat buffer."""

    def __init__(self, weight: torch.Tensor, bias: torch.Tensor, kind: str):
        super().__init__()
        self.kind = kind
        self.weight = nn.Parameter(weight)
        self.bias = nn.Parameter(bias)

    def extra_repr(self) -> str:
        return f"{self.kind}, weight={tuple(self.weight.shape)}"


class Network(nn.Module):
    """LeNet-5-style CNN: conv(3,6,5)-pool-conv(6,16,5)-pool-fc120-fc84-fc10."""

    def __init__(self):
        super().__init__()
        flat = torch.zeros(flat_layout.FLAT_NUMEL, dtype=torch.float32)
        grad = torch.zeros(flat_layout.FLAT_NUMEL, dtype=torch.float32)
        # not in state_dict: the checkpoint format stays the reference's 10 keys
        self.register_buffer("_flat", flat, persistent=False)
        self.register_buffer("_flat_grad", grad, persistent=False)
        v = flat_layout.views(self._flat)
        self.conv1 = _Params(v["conv1.weight"], v["conv1.bias"], "Conv2d(3, 6, 5)")
        self.pool = nn.MaxPool2d(2, 2)  # parameter-free; kept for attribute parity
        self.conv2 = _Params(v["conv2.weight"], v["conv2.bias"], "Conv2d(6, 16, 5)")
        self.fc1 = _Params(v["fc1.weight"], v["fc1.bias"], "Linear(400, 120)")
        self.fc2 = _Params(v["fc2.weight"], v["fc2.bias"], "Linear(120, 84)")
        self.fc3 = _Params(v["fc3.weight"], v["fc3.bias"], "Linear(84, 10)")
        self.reset_parameters()
        self._bind_grads()

    # ---- init: PyTorch's default for Conv2d/Linear, kaiming_uniform(a=sqrt(5))
    # which is U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias alike.
    @torch.no_grad()
    def reset_parameters(self) -> None:
        for m in (self.conv1, self.conv2, self.fc1, self.fc2, self.fc3):
            fan_in = m.weight[0].numel()
            bound = 1.0 / math.sqrt(fan_in)
            m.weight.uniform_(-bound, bound)
            m.bias.uniform_(-bound, bound)

    # ---- flat-buffer plumbing -------------------------------------------------
    def named_flat_params(self):
        for name in flat_layout.PARAM_SHAPES:
            mod, leaf = name.split(".")
            yield name, getattr(getattr(self, mod), leaf)

    def _rebind(self) -> None:
        """Re-point every Parameter at the (possibly moved) flat buffers."""
        v = flat_layout.views(self._flat)
        for name, p in self.named_flat_params():
            p.data = v[name]
        self._bind_grads()

    def _bind_grads(self) -> None:
        g = flat_layout.views(self._flat_grad)
        for name, p in self.named_flat_params():
            p.grad = g[name]

    def _apply(self, fn, recurse=True):
        # Move/cast the two flat buffers only, then re-view: nn.Module._apply on
        # the Parameters themselves would give each one private storage.
        new_flat = fn(self._flat)
        new_grad = fn(self._flat_grad)
        if new_flat.dtype != torch.float32:
            raise TypeError("Network keeps fp32 master parameters; cast inputs, not the model")
        self._buffers["_flat"] = new_flat.contiguous()
        self._buffers["_flat_grad"] = new_grad.contiguous()
        self._rebind()
        return self

    def __deepcopy__(self, memo):
        # nn.Parameter.__deepcopy__ clones each view into private storage, which would silently
        # detach the copy's parameters from its flat buffer; rebuild the aliasing instead.
        new = type(self)()
        new._buffers["_flat"] = self._flat.detach().clone()
        new._buffers["_flat_grad"] = self._flat_grad.detach().clone()
        new._rebind()
        new.train(self.training)
        memo[id(self)] = new
        return new

    @property
    def flat_params(self) -> torch.Tensor:
        return self._flat

    @property
    def flat_grads(self) -> torch.Tensor:
        return self._flat_grad

    def zero_grad(self, set_to_none: bool = False) -> None:  # noqa: D401
        """Zero the flat gradient buffer; gradients stay views of it."""
        self._flat_grad.zero_()
        self._bind_grads()

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        out = super().load_state_dict(state_dict, strict=strict, assign=False)
        return out

    # ---- forward ----------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = F_native.conv_relu_pool(x, self.conv1.weight, self.conv1.bias)
        x = F_native.conv_relu_pool(x, self.conv2.weight, self.conv2.bias)
        x = torch.flatten(x, 1)  # (C,H,W) order, as fc1.weight expects
        x = F_native.linear(x, self.fc1.weight, self.fc1.bias, relu=True)
        x = F_native.linear(x, self.fc2.weight, self.fc2.bias, relu=True)
        x = F_native.linear(x, self.fc3.weight, self.fc3.bias, relu=False)
        return x
