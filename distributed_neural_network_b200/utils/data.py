"""Datasets: CIFAR-10 (offline-safe) and a CIFAR-shaped synthetic generator.

Behavioural parity with the reference's data layer:
  * transform: ToTensor + Normalize(0.5, 0.5) per channel -> fp32 NCHW in [-1, 1]
    (reference: data_parallelism_train.py:24-27)
  * partition_dataset: contiguous shards of ``len // (size-1)`` for ranks 1..size-1,
    remainder dropped (reference: data_parallelism_train.py:49-53)
  * loaders: shuffled train batches, sequential test batches

B200-first differences: the whole dataset is materialised ONCE as two tensors
(50 000 x 3 x 32 x 32 fp32 = 614 MB, trivially resident in 180 GB of HBM or in pinned
host memory), batches are index views, and no DataLoader worker processes, pickling or
per-sample transforms sit on the step path. There is no network on the build or the GPU
box, so ``download=True`` is never attempted: real CIFAR-10 is used when
``<root>/cifar-10-batches-py`` exists (or the binary version ``cifar-10-batches-bin``, or the
undecompressed ``cifar-10-python.tar.gz``), otherwise a deterministic synthetic set of the
same shape stands in (and the caller is told).
"""
from __future__ import annotations

import os
import pickle
from dataclasses import dataclass
from typing import Iterator, List, Optional, Tuple

import numpy as np
import torch

CIFAR_DIR = "cifar-10-batches-py"
CIFAR_BIN_DIR = "cifar-10-batches-bin"          # the dataset's "binary version"
CIFAR_TARBALL = "cifar-10-python.tar.gz"        # what the reference's download leaves in ./data
TRAIN_SIZE, TEST_SIZE = 50_000, 10_000


@dataclass
class TensorDataset:
    """A whole dataset as two tensors (images fp32 NCHW in [-1,1], labels int64)."""

    images: torch.Tensor
    labels: torch.Tensor
    source: str = "synthetic"

    def __len__(self) -> int:
        return int(self.labels.shape[0])

    def __getitem__(self, i):
        return self.images[i], self.labels[i]

    def to(self, device, non_blocking: bool = False) -> "TensorDataset":
        return TensorDataset(self.images.to(device, non_blocking=non_blocking),
                             self.labels.to(device, non_blocking=non_blocking), self.source)

    def pin_memory(self) -> "TensorDataset":
        if not torch.cuda.is_available():
            return self
        return TensorDataset(self.images.pin_memory(), self.labels.pin_memory(), self.source)


class Subset:
    """Index view of a dataset; same contract as ``torch.utils.data.Subset`` (``.indices``, ``.dataset``)."""

    def __init__(self, dataset, indices):
        self.dataset = dataset
        self.indices = list(indices)

    def __len__(self) -> int:
        return len(self.indices)

    def __getitem__(self, i):
        return self.dataset[self.indices[i]]

    def materialize(self) -> TensorDataset:
        ds = self.dataset
        if isinstance(ds, Subset):
            ds = ds.materialize()
        if len(self.indices) > 0 and self.indices == list(range(self.indices[0], self.indices[0] + len(self.indices))):
            lo, hi = self.indices[0], self.indices[0] + len(self.indices)  # contiguous shard: a view, no copy
            return TensorDataset(ds.images[lo:hi], ds.labels[lo:hi], ds.source)
        idx = torch.as_tensor(self.indices, dtype=torch.long)
        return TensorDataset(ds.images[idx], ds.labels[idx], ds.source)


def transform(u8_nhwc: np.ndarray) -> torch.Tensor:
    """ToTensor + Normalize((.5,.5,.5),(.5,.5,.5)): uint8 NHWC -> fp32 NCHW in [-1, 1]."""
    t = torch.from_numpy(np.ascontiguousarray(u8_nhwc)).permute(0, 3, 1, 2).contiguous().float()
    return t.div_(255.0).sub_(0.5).div_(0.5)


def partition_dataset(dataset, rank: int, size: int) -> Subset:
    """Contiguous shard for child ``rank`` of ``size`` processes (rank 0 is the parent).

    Index-exact with reference data_parallelism_train.py:49-53.
    """
    total_size = len(dataset)
    partition_size = total_size // (size - 1)
    indices = list(range((rank - 1) * partition_size, rank * partition_size))
    return Subset(dataset, indices)


def partition_symmetric(dataset, rank: int, world: int) -> Subset:
    """All ``world`` ranks train: contiguous shards of ``len // world`` (remainder dropped)."""
    n = len(dataset) // world
    return Subset(dataset, list(range(rank * n, (rank + 1) * n)))


def synthetic(n: int, seed: int = 0, device: str = "cpu", learnable: bool = True) -> TensorDataset:
    """Deterministic CIFAR-10-shaped data: ``n`` images 3x32x32 in [-1,1], 10 classes.

    With ``learnable`` each class is a fixed random template plus noise, so loss curves
    behave like a real (easy) dataset; otherwise pure noise with random labels.
    """
    g = torch.Generator(device="cpu").manual_seed(seed)
    labels = torch.randint(0, 10, (n,), generator=g)
    noise = torch.rand((n, 3, 32, 32), generator=g) * 2.0 - 1.0
    if learnable:
        gt = torch.Generator(device="cpu").manual_seed(1234)  # templates shared by every split
        templates = torch.rand((10, 3, 32, 32), generator=gt) * 2.0 - 1.0
        images = (0.5 * templates[labels] + 0.5 * noise).clamp_(-1.0, 1.0)
    else:
        images = noise
    ds = TensorDataset(images.contiguous(), labels.contiguous(), "synthetic")
    return ds.to(device) if device != "cpu" else ds


def _load_cifar_batches(root: str, train: bool) -> Optional[Tuple[np.ndarray, np.ndarray]]:
    d = os.path.join(root, CIFAR_DIR)
    names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
    if not all(os.path.exists(os.path.join(d, n)) for n in names):
        return None
    xs: List[np.ndarray] = []
    ys: List[int] = []
    for n in names:
        with open(os.path.join(d, n), "rb") as f:
            entry = pickle.load(f, encoding="latin1")
        xs.append(entry["data"])
        ys.extend(entry.get("labels", entry.get("fine_labels")))
    x = np.vstack(xs).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)  # NHWC uint8, as torchvision holds it
    return x, np.asarray(ys, dtype=np.int64)


def _load_cifar_binary(root: str, train: bool) -> Optional[Tuple[np.ndarray, np.ndarray]]:
    """The "binary version" of the dataset (``cifar-10-batches-bin``): records of 1 label byte + 3072 bytes CHW."""
    d = os.path.join(root, CIFAR_BIN_DIR)
    names = [f"data_batch_{i}.bin" for i in range(1, 6)] if train else ["test_batch.bin"]
    if not all(os.path.exists(os.path.join(d, n)) for n in names):
        return None
    recs = [np.fromfile(os.path.join(d, n), dtype=np.uint8) for n in names]
    for n, r in zip(names, recs):
        if r.size == 0 or r.size % 3073 != 0:
            raise ValueError(f"{os.path.join(d, n)}: size {r.size} is not a multiple of the 3073-byte record")
    rec = np.concatenate(recs).reshape(-1, 3073)
    x = rec[:, 1:].reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)     # NHWC uint8, as the pickled batches are held
    return np.ascontiguousarray(x), rec[:, 0].astype(np.int64)


def _load_cifar_tarball(root: str, train: bool) -> Optional[Tuple[np.ndarray, np.ndarray]]:
    """``cifar-10-python.tar.gz`` as torchvision leaves it next to the extracted directory: read in place."""
    path = os.path.join(root, CIFAR_TARBALL)
    if not os.path.exists(path):
        return None
    import tarfile

    names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
    xs: List[np.ndarray] = []
    ys: List[int] = []
    with tarfile.open(path, "r:*") as tar:
        members = {os.path.basename(m.name): m for m in tar.getmembers() if m.isfile()}
        if not all(n in members for n in names):
            return None
        for n in names:
            entry = pickle.load(tar.extractfile(members[n]), encoding="latin1")
            xs.append(entry["data"])
            ys.extend(entry.get("labels", entry.get("fine_labels")))
    x = np.vstack(xs).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)
    return x, np.asarray(ys, dtype=np.int64)


def cifar10(root: str = "./data", train: bool = True, allow_synthetic: bool = True,
            synthetic_size: Optional[int] = None, verbose: bool = True) -> TensorDataset:
    """CIFAR-10 as tensors; synthetic stand-in of the same shape when the files are absent."""
    got = _load_cifar_batches(root, train) or _load_cifar_binary(root, train) or _load_cifar_tarball(root, train)
    if got is not None:
        x, y = got
        return TensorDataset(transform(x), torch.from_numpy(y), "cifar10")
    if not allow_synthetic:
        raise FileNotFoundError(
            f"{os.path.join(root, CIFAR_DIR)} not found and downloading is impossible here (no network)")
    n = synthetic_size if synthetic_size is not None else (TRAIN_SIZE if train else TEST_SIZE)
    if verbose:
        print(f"[data] CIFAR-10 files not found under {root!r}; using {n} synthetic CIFAR-shaped images")
    return synthetic(n, seed=0 if train else 1)


def batches(ds: TensorDataset, batch_size: int, shuffle: bool, generator: Optional[torch.Generator] = None,
            drop_last: bool = False) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
    """Mini-batches of a tensor dataset (DataLoader semantics: last short batch kept by default)."""
    n = len(ds)
    if shuffle:
        perm = torch.randperm(n, generator=generator)
        perm = perm.to(ds.images.device)
        for i in range(0, n, batch_size):
            idx = perm[i:i + batch_size]
            if drop_last and idx.numel() < batch_size:
                return
            yield ds.images[idx], ds.labels[idx]
    else:
        for i in range(0, n, batch_size):
            if drop_last and i + batch_size > n:
                return
            yield ds.images[i:i + batch_size], ds.labels[i:i + batch_size]


def epoch_tensor(ds: TensorDataset, batch_size: int, shuffle: bool, generator: Optional[torch.Generator] = None):
    """One epoch as ``x [steps, B, 3, 32, 32]``, ``y [steps, B]`` (the full batches; the layout the
    native multi-step loops consume) plus the trailing short batch ``(x_tail, y_tail)`` or ``None`` --
    DataLoader(drop_last=False) semantics."""
    n = len(ds)
    steps = n // batch_size
    perm = torch.randperm(n, generator=generator) if shuffle else torch.arange(n)
    perm = perm.to(ds.images.device)
    full = perm[: steps * batch_size]
    x = ds.images[full].view(steps, batch_size, 3, 32, 32)
    y = ds.labels[full].view(steps, batch_size)
    tail = None
    if n > steps * batch_size:
        rest = perm[steps * batch_size:]
        tail = (ds.images[rest].contiguous(), ds.labels[rest].contiguous())
    return x, y, tail
