"""Per-bucket gradient averaging on a side stream, overlapped with backward (layer-at-a-time path).

The fused step engine exchanges the whole flat gradient inside its update kernel; there is nothing left to
overlap it with. The autograd path (``Network.forward`` / ``loss.backward()`` through the layer kernels, the path
for batches of hundreds or thousands of images per GPU) is different: backward is a sequence of kernels, and the
gradient of the last layers is final long before the convolutions' is. Here the flat gradient is cut into three
buckets in the order backward finishes them,

    fc2 + fc3 (11 072 floats)  ->  fc1 (48 128 floats, 77 % of the bytes)  ->  conv1 + conv2 (2 944 floats)

and as soon as autograd has accumulated the last tensor of a bucket, the peer-memory all-reduce of exactly that slice
of ``Network.flat_grads`` (csrc/allreduce.cu, mean, in place) is enqueued on a side stream that waits for an event on
the backward stream. The bulk of the exchange therefore runs under the convolution backward kernels; only the small
conv bucket is exposed. ``finish()`` makes the main stream wait for the side stream before the optimizer kernel.

Replaces the reference's end-of-epoch send/recv of whole state_dicts (data_parallelism_train.py:210,226-244) for this
path; without a CUDA engine (CPU runs, tests) the same buckets go through the control plane's all-reduce, in the hook.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch

from ..ops import flat as L

# buckets as [lo, hi) ranges of the flat vector, in the order backward completes them
BUCKETS: List[Tuple[str, int, int, Tuple[str, ...]]] = [
    ("fc2+fc3", L.PARAM_OFFSETS["fc2.weight"], L.FLAT_NUMEL, ("fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias")),
    ("fc1", L.PARAM_OFFSETS["fc1.weight"], L.PARAM_OFFSETS["fc2.weight"], ("fc1.weight", "fc1.bias")),
    ("conv", 0, L.PARAM_OFFSETS["fc1.weight"], ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias")),
]


class BucketedGradientAverager:
    """Averages ``model.flat_grads`` over the ranks bucket by bucket while backward is still running.

    ``engine`` (``ops.native.Engine`` with a peer group attached) does the exchange on the GPU; without it
    ``allreduce_fn(tensor)`` -- in-place SUM over the ranks, e.g. ``comm.allreduce_sum_`` -- is used and the mean is
    taken here. ``overlap=False`` enqueues the same three all-reduces after backward instead (the A/B switch).
    """

    def __init__(self, model, world: int, engine=None, allreduce_fn: Optional[Callable] = None, variant="one-shot",
                 overlap: bool = True):
        self.model, self.world, self.engine, self.allreduce_fn = model, world, engine, allreduce_fn
        self.variant, self.overlap = variant, overlap
        self.launched: List[str] = []          # bucket names in the order they were exchanged (for tests / traces)
        self._pending = {}
        self._handles = []
        self._side = torch.cuda.Stream(device=model.flat_params.device) if engine is not None else None
        self._armed = False
        if world > 1:
            named = dict(model.named_flat_params())
            for bi, (_, _, _, names) in enumerate(BUCKETS):
                for n in names:
                    self._handles.append(named[n].register_post_accumulate_grad_hook(self._make_hook(bi, n)))

    def _make_hook(self, bi: int, name: str):
        def hook(_param):
            if not self._armed:
                return
            left = self._pending[bi]
            left.discard(name)
            if not left and self.overlap:
                self._exchange(bi)
        return hook

    def arm(self) -> None:
        """Call before ``backward()``: every bucket waits for all of its tensors again."""
        self._pending = {bi: set(names) for bi, (_, _, _, names) in enumerate(BUCKETS)}
        self.launched = []
        self._armed = self.world > 1

    def _exchange(self, bi: int) -> None:
        name, lo, hi, _ = BUCKETS[bi]
        g = self.model.flat_grads[lo:hi]
        self.launched.append(name)
        if self.engine is not None:
            main = torch.cuda.current_stream()
            if self.overlap:
                ev = torch.cuda.Event()
                ev.record(main)                       # this bucket's gradient kernels are in front of the event
                self._side.wait_event(ev)
                with torch.cuda.stream(self._side):
                    self.engine.allreduce(g, g, mode=0, two_shot=self.variant)
            else:
                self.engine.allreduce(g, g, mode=0, two_shot=self.variant)
        else:
            self.allreduce_fn(g)
            g.div_(self.world)

    def finish(self) -> None:
        """Call after ``backward()``: exchanges what has not gone out yet and orders the optimizer behind all of it."""
        if not self._armed:
            return
        for bi in range(len(BUCKETS)):
            if BUCKETS[bi][0] not in self.launched:
                self._exchange(bi)
        if self.engine is not None and self.overlap:
            torch.cuda.current_stream().wait_stream(self._side)
        self._armed = False

    def close(self) -> None:
        for h in self._handles:
            h.remove()
        self._handles = []
