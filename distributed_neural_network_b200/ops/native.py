"""ctypes binding of the in-tree native runtime (``_native/libdnnb200.so``).

The library is built by ``build.py`` / ``__graft_entry__.build()`` with
``nvcc -gencode arch=compute_100a,code=sm_100a`` and exposes the C ABI declared
in ``include/dnnb200.h``. Tensors are passed as raw device pointers and the
kernels are enqueued on PyTorch's current CUDA stream, so they compose with
torch ops, events and CUDA graphs.

What it stands in for in the reference: ``Engine.step / train_device / train_host`` are the body of
``run_child`` (data_parallelism_train.py:185-213: forward, loss, backward, SGD step per batch) and, with a peer
group attached, also ``run_parent``'s parameter exchange (data_parallelism_train.py:219-254) folded into every
step; ``Engine.evaluate`` is ``eval`` (data_parallelism_train.py:157-183); ``Engine.allreduce`` is the once-per-epoch
state_dict average of the parity modes (data_parallelism_train.py:226-244).

There is deliberately no fallback here: on a CUDA tensor a missing library is
an error (``require()``), so a GPU run can never silently measure eager PyTorch.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional, Sequence

import torch

_LIB_PATH = Path(os.environ.get("DNNB200_LIB") or
                 Path(__file__).resolve().parent.parent / "_native" / "libdnnb200.so")
_lib: Optional[C.CDLL] = None
_load_error: Optional[str] = None

FLAT_NUMEL = 62144
NVLS_FLOATS = 2 * FLAT_NUMEL + 2 * 16 * 128
IPC_HANDLE_BYTES = 64


class Ctl(C.Structure):
    """Mirror of ``dnnb200_ctl`` / ``StepCtl`` (csrc/step_types.h)."""

    _fields_ = [
        ("step", C.c_ulonglong),
        ("timeout_ns", C.c_ulonglong),
        ("lr", C.c_float),
        ("mu", C.c_float),
        ("loss_sum", C.c_float),
        ("loss_steps", C.c_uint),
        ("ctas_done", C.c_uint),
        ("status", C.c_uint),
        ("active_mask", C.c_uint),
        ("two_shot", C.c_uint),
        ("updates_done", C.c_uint),
        ("divisor_override", C.c_uint),
    ]


# Arithmetic of the conv stages of the clustered step kernel: "fp32" = FFMA throughout (matches PyTorch fp32 to 2e-5),
# "tf32" = tcgen05 implicit GEMMs (TF32 operands, fp32 accumulate). Both are built and tested; the default is the one
# that MEASURES faster at the benchmark's batch size -- fp32, by 1.3 us per step (profiles/step_tc_ab_v1.md: the
# model's GEMMs are so small that a step needs ~50 dependent tcgen05.mma of M=128 N=16 K=8, each of which costs the
# issuing thread 64 cycles whatever its size).
_default_precision = os.environ.get("DNNB200_PRECISION", "fp32")


PRECISIONS = {"fp32": 0, "tf32": 1, "tf32-dgrad": 2}


def set_default_precision(p: str) -> None:
    global _default_precision
    if p not in PRECISIONS:
        raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
    _default_precision = p


def default_precision() -> str:
    return _default_precision


EXCHANGE_VARIANTS = {"one-shot": 0, "two-shot": 1, "nvls": 2, "one-shot-128": 3}


def variant_id(v) -> int:
    """Exchange variant as the C ABI numbers it (include/dnnb200.h); accepts the name, the number or a bool."""
    if isinstance(v, str):
        return EXCHANGE_VARIANTS[v]
    return int(v)


def lib_path() -> Path:
    return _LIB_PATH


def _load() -> Optional[C.CDLL]:
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    if not _LIB_PATH.exists():
        _load_error = f"{_LIB_PATH} not found -- run `python build.py`"
        return None
    try:
        lib = C.CDLL(str(_LIB_PATH))
    except OSError as e:  # e.g. libcudart missing on a CPU-only box
        _load_error = f"cannot load {_LIB_PATH}: {e}"
        return None
    lib.dnnb200_last_error.restype = C.c_char_p
    lib.dnnb200_engine_launches.restype = C.c_ulonglong
    lib.dnnb200_engine_launches.argtypes = [C.c_void_p]
    _lib = lib
    return lib


def available() -> bool:
    return _load() is not None


def require() -> C.CDLL:
    lib = _load()
    if lib is None:
        raise RuntimeError(
            "native sm_100a runtime is not available: " + str(_load_error) +
            " (CUDA tensors are never routed to an eager fallback)")
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        msg = require().dnnb200_last_error().decode(errors="replace")
        raise RuntimeError(f"dnnb200 native call failed (code {rc}): {msg}")


def _p(t: Optional[torch.Tensor]) -> C.c_void_p:
    if t is None:
        return C.c_void_p(0)
    return C.c_void_p(t.data_ptr())


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda:
        raise TypeError(f"{name}: expected a contiguous CUDA float32 tensor, got {t.dtype} "
                        f"contiguous={t.is_contiguous()} device={t.device}")
    return t


def layout():
    lib = require()
    offs = (C.c_int * 10)()
    flat, sl, ns = C.c_int(), C.c_int(), C.c_int()
    _check(lib.dnnb200_layout(offs, C.byref(flat), C.byref(sl), C.byref(ns)))
    return list(offs), flat.value, sl.value, ns.value


# ---------------------------------------------------------------------------
# layer kernels
# ---------------------------------------------------------------------------
def conv_relu_pool_fwd(x, w, b, y, idx) -> None:
    lib = require()
    N, Ci, H, W = x.shape
    _check(lib.dnnb200_conv_relu_pool_fwd(_p(_f32(x, "x")), _p(_f32(w, "w")), _p(_f32(b, "b")), _p(_f32(y, "y")),
                                          _p(idx), N, Ci, H, W, w.shape[0], _stream()))


def tc_conv_relu_pool_fwd(x, w, b, y, idx) -> None:
    """Tensor-core (tcgen05, TF32) version of conv_relu_pool_fwd for the CNN's two conv shapes."""
    lib = require()
    N, Ci, H, W = x.shape
    _check(lib.dnnb200_tc_conv_relu_pool_fwd(_p(_f32(x, "x")), _p(_f32(w, "w")), _p(_f32(b, "b")), _p(_f32(y, "y")),
                                             _p(idx), N, Ci, H, W, w.shape[0], _stream()))


def tc_conv_supported(x, w) -> bool:
    return tuple(x.shape[1:]) + (w.shape[0],) in ((3, 32, 32, 6), (6, 14, 14, 16))


def tc_gemm_probe(A, B, D) -> None:
    _check(require().dnnb200_tc_gemm_probe(_p(_f32(A, "A")), _p(_f32(B, "B")), _p(_f32(D, "D")), A.shape[1], _stream()))


def conv_relu_pool_bwd(x, w, y, idx, dy, dx, dw, db, tensor_cores: bool = False) -> None:
    lib = require()
    N, Ci, H, W = x.shape
    fn = lib.dnnb200_conv_relu_pool_bwd_tc if tensor_cores else lib.dnnb200_conv_relu_pool_bwd
    _check(fn(_p(_f32(x, "x")), _p(_f32(w, "w")), _p(_f32(y, "y")), _p(idx),
                                          _p(_f32(dy, "dy")), _p(dx), _p(_f32(dw, "dw")), _p(_f32(db, "db")),
                                          N, Ci, H, W, w.shape[0], _stream()))


def tc_conv_wgrad(x, y, idx, dy, dw, db) -> None:
    """Tensor-core (tcgen05, TF32) weight/bias gradient of conv_relu_pool for the CNN's two conv shapes
    (csrc/tc_conv_wgrad.cu). The measured alternative to the pooled-grid FFMA kernel inside conv_relu_pool_bwd."""
    lib = require()
    N, Ci, H, W = x.shape
    _check(lib.dnnb200_tc_conv_wgrad(_p(_f32(x, "x")), _p(_f32(y, "y")), _p(idx), _p(_f32(dy, "dy")), _p(_f32(dw, "dw")),
                                     _p(_f32(db, "db")), N, Ci, H, W, dw.shape[0], _stream()))


def linear_fwd(x, w, b, y, relu: bool) -> None:
    lib = require()
    N, K = x.shape
    _check(lib.dnnb200_linear_fwd(_p(_f32(x, "x")), _p(_f32(w, "w")), _p(_f32(b, "b")), _p(_f32(y, "y")),
                                  N, w.shape[0], K, int(relu), _stream()))


def tc_linear_fwd(x, w, b, y, relu: bool) -> None:
    """Tensor-core (tcgen05 + TMA, TF32) version of linear_fwd; needs M <= 128 and K % 4 == 0."""
    lib = require()
    N, K = x.shape
    _check(lib.dnnb200_tc_linear_fwd(_p(_f32(x, "x")), _p(_f32(w, "w")), _p(_f32(b, "b")), _p(_f32(y, "y")),
                                     N, w.shape[0], K, int(relu), _stream()))


def tc_linear_supported(x, w) -> bool:
    return w.shape[0] <= 128 and x.shape[1] % 4 == 0 and x.data_ptr() % 16 == 0 and w.data_ptr() % 16 == 0


def tc_lab(a_img, b_img, adesc: int, bdesc: int, idesc: int, ncols: int, nmma: int = 1, a_step: int = 0,
           b_step: int = 0):
    """Descriptor laboratory: raw smem images + raw descriptors -> D[128, ncols] (csrc/tc_lab.cu)."""
    lib = require()
    D = torch.full((128, ncols), float("nan"), device=a_img.device)
    _check(lib.dnnb200_tc_lab(_p(_f32(a_img, "a_img")), a_img.numel(), _p(_f32(b_img, "b_img")), b_img.numel(),
                              C.c_ulonglong(adesc), C.c_ulonglong(bdesc), C.c_uint(idesc), nmma, a_step, b_step,
                              _p(D), ncols, _stream()))
    return D


def tc_linear_bwd_supported(x, w, dy) -> bool:
    M = w.shape[0]
    return (tc_linear_supported(x, w) and M % 4 == 0 and M >= 4 and dy.data_ptr() % 16 == 0)


def tc_linear_bwd(x, w, y, dy, dx, dw, db, relu: bool) -> None:
    """Tensor-core (tcgen05 + TMA, TF32) linear backward: dX, dW and db (csrc/tc_linear_bwd.cu)."""
    lib = require()
    N, K = x.shape
    scratch = torch.empty_like(dy) if relu else None
    _check(lib.dnnb200_tc_linear_bwd(_p(_f32(x, "x")), _p(_f32(w, "w")), _p(_f32(y, "y")), _p(_f32(dy, "dy")),
                                     _p(scratch), _p(dx), _p(_f32(dw, "dw")), _p(_f32(db, "db")),
                                     N, w.shape[0], K, int(relu), _stream()))


def linear_bwd(x, w, y, dy, dx, dw, db, relu: bool) -> None:
    lib = require()
    N, K = x.shape
    _check(lib.dnnb200_linear_bwd(_p(_f32(x, "x")), _p(_f32(w, "w")), _p(_f32(y, "y")), _p(_f32(dy, "dy")), _p(dx),
                                  _p(_f32(dw, "dw")), _p(_f32(db, "db")), N, w.shape[0], K, int(relu), _stream()))


def softmax_xent(logits, target, loss, dlogits) -> None:
    lib = require()
    if target.dtype != torch.int64:
        raise TypeError("target must be int64")
    N, Cn = logits.shape
    _check(lib.dnnb200_softmax_xent(_p(_f32(logits, "logits")), _p(target), _p(loss), _p(_f32(dlogits, "dlogits")),
                                    N, Cn, _stream()))


def sgd_momentum(p, g, buf, lr: float, mu: float, first_step: bool, scale: float = 1.0) -> None:
    lib = require()
    _check(lib.dnnb200_sgd_momentum(_p(_f32(p, "p")), _p(_f32(g, "g")), _p(_f32(buf, "buf")), p.numel(),
                                    C.c_float(lr), C.c_float(mu), int(first_step), C.c_float(scale), _stream()))


# ---------------------------------------------------------------------------
# peer group
# ---------------------------------------------------------------------------
def share_fd(rank: int, world: int, name: str, word: int = 0, fd: int = -1, timeout_s: float = 30.0):
    """Rank 0 hands (word, file descriptor) to every other rank on this host over an abstract Unix socket named after
    ``name`` (SCM_RIGHTS); returns (word, fd) -- on ranks > 0 a duplicate of the descriptor, theirs to close."""
    w, f = C.c_ulonglong(word), C.c_int(fd)
    _check(require().dnnb200_share_fd(rank, world, name.encode(), C.byref(w), C.byref(f), int(timeout_s * 1000)))
    return int(w.value), int(f.value)


def nvls_supported(device_index: int) -> bool:
    lib = _load()
    return bool(lib is not None and torch.cuda.is_available() and lib.dnnb200_nvls_supported(device_index))


class NvlsGroup:
    """NVLink multicast group (csrc/nvls_group.cu): ``NvlsGroup(...)`` on every rank, a barrier, ``map()`` on every
    rank, a barrier, then ``local_ptr`` / ``multicast_ptr`` are this rank's and the all-GPU view of one zero-filled
    symmetric buffer. ``name`` must be the same on all ranks and unique to the job."""

    def __init__(self, rank: int, world: int, device_index: int, nbytes: int, name: str, timeout_s: float = 30.0):
        self._h = C.c_void_p()
        self.local_ptr = self.multicast_ptr = 0
        self.nbytes = 0
        _check(require().dnnb200_nvls_open(C.byref(self._h), rank, world, device_index, C.c_size_t(nbytes),
                                           name.encode(), int(timeout_s * 1000)))

    def map(self):
        loc, mc, nb = C.c_void_p(), C.c_void_p(), C.c_size_t()
        _check(require().dnnb200_nvls_map(self._h, C.byref(loc), C.byref(mc), C.byref(nb)))
        self.local_ptr, self.multicast_ptr, self.nbytes = int(loc.value), int(mc.value), int(nb.value)
        return self.local_ptr, self.multicast_ptr

    def close(self) -> None:
        if self._h:
            require().dnnb200_nvls_close(self._h)
            self._h = C.c_void_p()
            self.local_ptr = self.multicast_ptr = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Group:
    """This rank's receive/flag area plus the mapped areas of its peers."""

    def __init__(self, rank: int, world: int):
        lib = require()
        self.rank, self.world = rank, world
        self._h = C.c_void_p()
        buf = (C.c_ubyte * IPC_HANDLE_BYTES)()
        _check(lib.dnnb200_group_create(rank, world, C.byref(self._h), buf))
        self.handle: bytes = bytes(buf)

    def connect(self, all_handles: Sequence[bytes]) -> None:
        assert len(all_handles) == self.world
        blob = b"".join(all_handles)
        assert len(blob) == self.world * IPC_HANDLE_BYTES
        _check(require().dnnb200_group_connect(self._h, blob))

    @staticmethod
    def connect_local(groups: Sequence["Group"]) -> None:
        arr = (C.c_void_p * len(groups))(*[g._h for g in groups])
        _check(require().dnnb200_group_connect_local(arr, len(groups)))

    def close(self) -> None:
        if self._h:
            require().dnnb200_group_destroy(self._h)
            self._h = C.c_void_p()


# ---------------------------------------------------------------------------
# fused step engine
# ---------------------------------------------------------------------------
class Engine:
    """Owns the step buffers; updates ``params``/``momentum`` (flat CUDA tensors) in place."""

    def __init__(self, params: torch.Tensor, momentum: torch.Tensor, grads: Optional[torch.Tensor],
                 max_batch: int, lr: float, mu: float, precision: Optional[str] = None):
        lib = require()
        for name, t in (("params", params), ("momentum", momentum)):
            _f32(t, name)
            if t.numel() != FLAT_NUMEL:
                raise ValueError(f"{name} must have {FLAT_NUMEL} elements")
        self.params, self.momentum, self.grads = params, momentum, grads  # keep alive
        self.max_batch = max_batch
        self._h = C.c_void_p()
        self.group: Optional[Group] = None
        with torch.cuda.device(params.device):
            _check(lib.dnnb200_engine_create(_p(params), _p(momentum), _p(grads), max_batch, C.c_float(lr),
                                             C.c_float(mu), C.byref(self._h)))
        self.precision = ""
        self.set_precision(precision or _default_precision)

    # -- configuration
    def attach_group(self, group: Optional[Group]) -> None:
        self.group = group
        _check(require().dnnb200_engine_attach_group(self._h, group._h if group else None))

    def set_hparams(self, lr: float, mu: float) -> None:
        _check(require().dnnb200_engine_set_hparams(self._h, C.c_float(lr), C.c_float(mu), _stream()))

    def set_membership(self, active_mask: int, timeout_ns: int = 0) -> None:
        _check(require().dnnb200_engine_set_membership(self._h, C.c_uint(active_mask), C.c_ulonglong(timeout_ns),
                                                       _stream()))

    def set_exchange_variant(self, variant) -> None:
        """Fused exchange: 0/False/"one-shot" (16-byte lines) and 3/"one-shot-128" (128-byte lines) tolerate ranks
        leaving; 1/True/"two-shot" and 2/"nvls" (in-switch reduction) need a static membership."""
        _check(require().dnnb200_engine_set_exchange_variant(self._h, variant_id(variant), _stream()))

    def set_divisor(self, divisor: int) -> None:
        """0 = divide exchanged sums by the ranks present; n = divide by n."""
        _check(require().dnnb200_engine_set_divisor(self._h, C.c_uint(divisor), _stream()))

    def set_precision(self, precision: str) -> None:
        """"fp32": FFMA throughout. "tf32": conv1/conv2 forward and the conv2 data gradient of the clustered step kernel
        run as implicit GEMMs on the tensor cores (tcgen05, TF32 operands, fp32 accumulation). "tf32-dgrad": only the
        conv2 data gradient does -- forward, loss and every gradient but conv1's stay bit-for-bit the fp32 ones."""
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        _check(require().dnnb200_engine_set_precision(self._h, PRECISIONS[precision]))
        self.precision = precision

    def zero_momentum(self) -> None:
        _check(require().dnnb200_engine_zero_momentum(self._h, _stream()))

    def reset_loss(self) -> None:
        _check(require().dnnb200_engine_reset_loss(self._h, _stream()))

    def read_ctl(self) -> Ctl:
        c = Ctl()
        _check(require().dnnb200_engine_read_ctl(self._h, C.byref(c), _stream()))
        return c

    def set_cluster_size(self, cs: int) -> None:
        """CTAs per image in the forward/backward kernel (0 = automatic, 1/2/4/8)."""
        _check(require().dnnb200_engine_set_cluster_size(self._h, cs))

    def set_trace(self, trace: Optional[torch.Tensor]) -> None:
        """int64 CUDA tensor [n_ctas*16] receiving per-stage SM clocks (development aid)."""
        self._trace = trace
        _check(require().dnnb200_engine_set_trace(self._h, _p(trace)))

    def pingpong(self, peer: int, iters: int = 1000) -> float:
        """One-way latency (us) of a 16-byte LL line to rank `peer`; the peer must call pingpong(my_rank) too."""
        out = C.c_double()
        _check(require().dnnb200_engine_pingpong(self._h, peer, iters, C.byref(out), _stream()))
        return float(out.value)

    def set_debug_flags(self, flags: int) -> None:
        _check(require().dnnb200_engine_set_debug_flags(self._h, flags))

    def set_update_trace(self, trace: Optional[torch.Tensor]) -> None:
        """int64 CUDA tensor [122*8] (or [1024*122*8] with debug flag 4): update-kernel timestamps per slice (development aid)."""
        self._utrace = trace
        _check(require().dnnb200_engine_set_update_trace(self._h, _p(trace)))

    def cluster_info(self, B: int):
        chosen = C.c_int()
        mc = (C.c_int * 9)()
        _check(require().dnnb200_engine_cluster_info(self._h, B, C.byref(chosen), mc))
        return chosen.value, {cs: mc[cs] for cs in range(2, 9)}

    @property
    def launches(self) -> int:
        return int(require().dnnb200_engine_launches(self._h))

    # -- work
    def step(self, x: torch.Tensor, y: torch.Tensor) -> None:
        _f32(x, "x")
        if y.dtype != torch.int64 or not y.is_cuda:
            raise TypeError("y must be a CUDA int64 tensor")
        _check(require().dnnb200_engine_step(self._h, _p(x), _p(y), x.shape[0], _stream()))

    def train_device(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """x [steps,B,3,32,32] fp32 and y [steps,B] int64 on the device: steps optimizer steps, one native loop."""
        _f32(x, "x")
        steps, B = y.shape
        assert x.shape[0] == steps and x.shape[1] == B and y.dtype == torch.int64 and y.is_contiguous()
        _check(require().dnnb200_engine_train_device(self._h, _p(x), _p(y), B, steps, _stream()))

    def grad_only(self, x: torch.Tensor, y: torch.Tensor, grads_out: torch.Tensor) -> float:
        _f32(x, "x"); _f32(grads_out, "grads_out")
        loss = C.c_float()
        _check(require().dnnb200_engine_grad_only(self._h, _p(x), _p(y), x.shape[0], _p(grads_out), C.byref(loss),
                                                  _stream()))
        return float(loss.value)

    def train_host(self, h_x: torch.Tensor, h_y: torch.Tensor, h_losses: Optional[torch.Tensor]) -> None:
        """h_x [steps,B,3,32,32] fp32 and h_y [steps,B] int64 in PINNED host memory."""
        if not (h_x.is_pinned() and h_y.is_pinned()):
            raise ValueError("train_host needs pinned host tensors")
        if h_losses is not None and not h_losses.is_pinned():
            raise ValueError("h_losses must be pinned")
        steps, B = h_y.shape
        assert h_x.shape[0] == steps and h_x.shape[1] == B and h_x.is_contiguous() and h_y.is_contiguous()
        _check(require().dnnb200_engine_train_host(self._h, _p(h_x), _p(h_y), B, steps, _p(h_losses), _stream()))

    def evaluate(self, x: torch.Tensor, y: torch.Tensor, bs: int, logits: Optional[torch.Tensor] = None):
        _f32(x, "x")
        out = (C.c_double * 2)()
        _check(require().dnnb200_engine_eval(self._h, _p(x), _p(y), x.shape[0], bs, _p(logits), out, _stream()))
        return float(out[0]), int(out[1])

    def allreduce(self, inp: torch.Tensor, out: Optional[torch.Tensor], mode: int = 0, two_shot=False) -> None:
        """``two_shot`` names the variant: False/0 one-shot, True/1 two-shot, 2/"nvls" in-switch (multimem),
        3/"one-shot-128" one-shot with 128-byte lines."""
        _f32(inp, "in")
        _check(require().dnnb200_engine_allreduce(self._h, _p(inp), _p(out), inp.numel(), mode, variant_id(two_shot),
                                                  _stream()))

    def allreduce_timed(self, inp: torch.Tensor, out: torch.Tensor, variant, iters: int = 200) -> float:
        """Device-timed microseconds per stand-alone mean all-reduce (a burst enqueued from C). Collective."""
        _f32(inp, "in")
        us = C.c_double()
        _check(require().dnnb200_engine_allreduce_timed(self._h, _p(inp), _p(out), inp.numel(), variant_id(variant),
                                                        iters, C.byref(us), _stream()))
        return float(us.value)

    def ll128_probe(self, peer: int, iters: int = 2000, n_ctas: int = 32):
        """Ping-pong of 128-byte lines with rank ``peer`` (which must call it too): returns (torn words -- must be 0,
        microseconds per round trip, gave_up)."""
        out = (C.c_ulonglong * 3)()
        _check(require().dnnb200_engine_ll128_probe(self._h, peer, iters, n_ctas, out, _stream()))
        return int(out[0]), float(out[1]) / 1e3 / iters, bool(out[2])

    def attach_nvls(self, local_buf, multicast_ptr: int) -> None:
        """``local_buf``: this rank's view of the symmetric buffer (an ``NvlsGroup`` after ``map()``, or a float32
        CUDA tensor of at least NVLS_FLOATS elements); ``multicast_ptr`` the all-GPU view of the same bytes."""
        if isinstance(local_buf, NvlsGroup):
            assert local_buf.nbytes >= NVLS_FLOATS * 4 and local_buf.local_ptr
            ptr = C.c_void_p(local_buf.local_ptr)
        else:
            _f32(local_buf, "local_buf")
            assert local_buf.numel() >= NVLS_FLOATS
            ptr = _p(local_buf)
        self._nvls_buf = local_buf              # keeps the mapping alive as long as the engine may use it
        _check(require().dnnb200_engine_attach_nvls(self._h, ptr, C.c_void_p(multicast_ptr)))

    def leave(self) -> None:
        _check(require().dnnb200_engine_leave(self._h, _stream()))

    def rejoin(self, step: int) -> None:
        _check(require().dnnb200_engine_rejoin(self._h, C.c_ulonglong(step), _stream()))

    def close(self) -> None:
        if self._h:
            require().dnnb200_engine_destroy(self._h)
            self._h = C.c_void_p()


def vr_step(engines: Sequence[Engine], xs: Sequence[Optional[torch.Tensor]], ys: Sequence[Optional[torch.Tensor]],
            B: int) -> None:
    """One step of ``len(engines)`` ranks emulated on ONE device (cooperative launch)."""
    n = len(engines)
    eh = (C.c_void_p * n)(*[e._h for e in engines])
    xp = (C.c_void_p * n)(*[(x.data_ptr() if x is not None else 0) for x in xs])
    yp = (C.c_void_p * n)(*[(y.data_ptr() if y is not None else 0) for y in ys])
    _check(require().dnnb200_vr_step(eh, n, xp, yp, B, _stream()))


def vr_allreduce(engines: Sequence[Engine], ins: Sequence[Optional[torch.Tensor]],
                 outs: Sequence[Optional[torch.Tensor]], count: int, mode: int = 0, two_shot: bool = False) -> None:
    n = len(engines)
    eh = (C.c_void_p * n)(*[e._h for e in engines])
    ip = (C.c_void_p * n)(*[(t.data_ptr() if t is not None else 0) for t in ins])
    op = (C.c_void_p * n)(*[(t.data_ptr() if t is not None else 0) for t in outs])
    _check(require().dnnb200_vr_allreduce(eh, n, ip, op, count, mode, variant_id(two_shot), _stream()))
