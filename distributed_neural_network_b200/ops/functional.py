"""Layer-level ops: one autograd Function per fused kernel, with a CPU oracle path.

On CUDA tensors every op runs a hand-written sm_100a kernel from the in-tree
native library (``ops/native.py``); if the library is missing on a GPU box the
call raises -- there is no silent eager fallback on the device. On CPU tensors
the same functions run plain PyTorch, which is the reference's own code path
(reference: models/model.py:20-27, data_parallelism_train.py:29) and serves as
the numerics oracle in tests.

Fusions: K1-K3/K4-K6 conv+bias+ReLU+maxpool in one kernel (argmax
kept as 2 bits per pooled pixel), K14/K15 ReLU/pool backward folded into the
conv backward prologue, K8-K10 linear+bias(+ReLU), K11/K12 softmax-CE forward and
backward in one pass.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import native


# Precision of the GEMM-shaped FORWARD ops (conv, linear) on CUDA: "tf32" runs the tcgen05 kernels (TF32
# operands, fp32 accumulate -- what PyTorch itself does for cuDNN convolutions by default), "fp32" the FFMA
# kernels. Backward and the fused step engine are fp32 either way.
_conv_precision = "tf32"


def set_conv_precision(mode: str) -> None:
    global _conv_precision
    if mode not in ("tf32", "fp32"):
        raise ValueError("conv precision must be 'tf32' or 'fp32'")
    _conv_precision = mode


def get_conv_precision() -> str:
    return _conv_precision


def _use_native(t: torch.Tensor) -> bool:
    if not t.is_cuda:
        return False
    native.require()  # raises with a clear message if the .so is not built
    return True


# ----------------------------------------------------------------------------
# conv5x5 (valid, stride 1) + bias + ReLU + maxpool 2x2/2
# ----------------------------------------------------------------------------
class _ConvReluPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        x = x.contiguous().float()
        N, Ci, H, W = x.shape
        Co = w.shape[0]
        Hp, Wp = (H - 4) // 2, (W - 4) // 2
        y = torch.empty((N, Co, Hp, Wp), device=x.device, dtype=torch.float32)
        idx = torch.empty((N, Co, Hp, Wp), device=x.device, dtype=torch.uint8)
        if _conv_precision == "tf32" and native.tc_conv_supported(x, w):
            native.tc_conv_relu_pool_fwd(x, w.contiguous(), b.contiguous(), y, idx)
        else:
            native.conv_relu_pool_fwd(x, w.contiguous(), b.contiguous(), y, idx)
        ctx.save_for_backward(x, w, y, idx)
        ctx.needs_dx = x.requires_grad
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y, idx = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(x) if ctx.needs_dx else None
        dw = torch.empty_like(w)
        db = torch.empty(w.shape[0], device=w.device, dtype=torch.float32)
        native.conv_relu_pool_bwd(x, w.contiguous(), y, idx, dy, dx, dw, db, tensor_cores=_conv_precision == "tf32")
        return dx, dw, db


def conv_relu_pool(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``max_pool2d(relu(conv2d(x, w, b)), 2, 2)`` (reference: models/model.py:21-22)."""
    if _use_native(x):
        return _ConvReluPool.apply(x, w, b)
    return F.max_pool2d(F.relu(F.conv2d(x, w, b)), 2, 2)


# ----------------------------------------------------------------------------
# linear + bias (+ ReLU)
# ----------------------------------------------------------------------------
# below this batch size the two scalar kernels of layers.cu are faster than prep + two tensor-core GEMMs
# (launch-bound: 33 vs 35 us at 256 rows of fc1, 36 vs 114 us at 1024; bench/tc_bwd_probe.py)
_TC_LINEAR_BWD_MIN_BATCH = 256


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, relu):
        x = x.contiguous().float()
        y = torch.empty((x.shape[0], w.shape[0]), device=x.device, dtype=torch.float32)
        w = w.contiguous()
        if _conv_precision == "tf32" and native.tc_linear_supported(x, w):
            native.tc_linear_fwd(x, w, b.contiguous(), y, relu)
        else:
            native.linear_fwd(x, w, b.contiguous(), y, relu)
        ctx.save_for_backward(x, w, y)
        ctx.relu = relu
        ctx.needs_dx = x.requires_grad
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(x) if ctx.needs_dx else None
        dw = torch.empty_like(w)
        db = torch.empty(w.shape[0], device=w.device, dtype=torch.float32)
        w = w.contiguous()
        if (_conv_precision == "tf32" and x.shape[0] >= _TC_LINEAR_BWD_MIN_BATCH
                and native.tc_linear_bwd_supported(x, w, dy)):
            native.tc_linear_bwd(x, w, y, dy, dx, dw, db, ctx.relu)
        else:
            native.linear_bwd(x, w, y, dy, dx, dw, db, ctx.relu)
        return dx, dw, db, None


def linear(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, relu: bool = False) -> torch.Tensor:
    """``relu?(x @ w.T + b)`` (reference: models/model.py:24-26)."""
    if _use_native(x):
        return _Linear.apply(x, w, b, relu)
    y = F.linear(x, w, b)
    return F.relu(y) if relu else y


# ----------------------------------------------------------------------------
# softmax cross-entropy, mean-reduced, int64 targets
# ----------------------------------------------------------------------------
class _SoftmaxXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        logits = logits.contiguous().float()
        loss = torch.empty((), device=logits.device, dtype=torch.float32)
        dlogits = torch.empty_like(logits)
        native.softmax_xent(logits, target.contiguous(), loss, dlogits)
        ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (dlogits,) = ctx.saved_tensors
        return dlogits * dloss, None


def cross_entropy(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """``nn.CrossEntropyLoss()(logits, target)`` (reference: data_parallelism_train.py:29)."""
    if _use_native(logits):
        return _SoftmaxXent.apply(logits, target)
    return F.cross_entropy(logits, target)


class CrossEntropyLoss(torch.nn.Module):
    """Drop-in for ``nn.CrossEntropyLoss()`` that uses the fused kernel on CUDA."""

    def forward(self, logits, target):
        return cross_entropy(logits, target)


# ----------------------------------------------------------------------------
# SGD with momentum on a flat buffer (dampening 0, no Nesterov, no weight decay)
# ----------------------------------------------------------------------------
@torch.no_grad()
def sgd_momentum_(p: torch.Tensor, g: torch.Tensor, buf: torch.Tensor, lr: float,
                  momentum: float, first_step: bool, grad_scale: float = 1.0) -> None:
    """In-place ``buf = mu*buf + g*scale`` (``buf = g*scale`` on the first step, as
    torch.optim.SGD does), then ``p -= lr*buf`` (reference: data_parallelism_train.py:187,199)."""
    if _use_native(p):
        native.sgd_momentum(p, g, buf, lr, momentum, first_step, grad_scale)
        return
    gs = g * grad_scale if grad_scale != 1.0 else g
    if first_step or momentum == 0.0:
        buf.copy_(gs)
    else:
        buf.mul_(momentum).add_(gs)
    p.add_(buf, alpha=-lr)
