"""Rank discovery and the control-plane communicator.

The reference bootstraps with ``MPI.COMM_WORLD`` / ``Get_rank`` / ``Get_size``
(reference: data_parallelism_train.py:60-62) and then moves whole state_dicts through
pickled blocking send/recv. Here the bootstrap is ONLY a control plane -- rank discovery,
the 64-byte CUDA-IPC handle exchange, barriers at epoch boundaries, tiny object broadcasts;
gradients never travel through it.

Two providers behind one interface:
  * mpi4py, when importable (the documented ``mpiexec -n P python ...`` command line)
  * environment + torch.distributed (gloo) otherwise -- what this image has. Recognised
    variables: OMPI_COMM_WORLD_RANK/SIZE, PMI_RANK/SIZE, PMIX_RANK, SLURM_PROCID/NTASKS,
    RANK/WORLD_SIZE (torchrun and the bundled launcher ``parallel/launch.py``).
"""
from __future__ import annotations

import datetime
import os
from typing import Any, List, Optional, Tuple

_RANK_VARS = ("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "PMIX_RANK", "SLURM_PROCID")
_SIZE_VARS = ("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS")
_LOCAL_VARS = ("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID")


def _first_env(names) -> Optional[int]:
    for n in names:
        v = os.environ.get(n)
        if v is not None and v != "":
            try:
                return int(v)
            except ValueError:
                pass
    return None


def discover() -> Tuple[int, int, int]:
    """(rank, world, local_rank) from the launcher's environment; (0, 1, 0) when run alone."""
    rank = _first_env(_RANK_VARS)
    world = _first_env(_SIZE_VARS)
    if rank is None or world is None:
        return 0, 1, 0
    local = _first_env(_LOCAL_VARS)
    return rank, world, (local if local is not None else rank)


class PeerLost(RuntimeError):
    """A control-plane collective could not complete: some rank is gone (process died, host unreachable)."""


class Comm:
    """Control-plane communicator: object all-gather / broadcast, barrier, CPU tensor all-reduce."""

    rank: int = 0
    world: int = 1
    local_rank: int = 0
    provider: str = "single"

    def barrier(self) -> None:
        pass

    def allgather(self, obj: Any) -> List[Any]:
        return [obj]

    def bcast(self, obj: Any, root: int = 0) -> Any:
        return obj

    def allreduce_sum_(self, t):  # CPU (or NCCL) tensor, in place
        return t

    def close(self) -> None:
        pass


class _MpiComm(Comm):
    provider = "mpi4py"

    def __init__(self):
        from mpi4py import MPI  # type: ignore

        self._MPI = MPI
        self._c = MPI.COMM_WORLD
        self.rank, self.world = self._c.Get_rank(), self._c.Get_size()
        self.local_rank = _first_env(_LOCAL_VARS) or 0

    def barrier(self):
        self._c.Barrier()

    def allgather(self, obj):
        return self._c.allgather(obj)

    def bcast(self, obj, root=0):
        return self._c.bcast(obj, root=root)

    def allreduce_sum_(self, t):
        import torch

        buf = t.detach().cpu().contiguous().numpy()
        out = buf.copy()
        self._c.Allreduce(buf, out, op=self._MPI.SUM)
        t.copy_(torch.from_numpy(out))
        return t


class _TorchComm(Comm):
    provider = "torch.distributed"

    def __init__(self, rank: int, world: int, local_rank: int, timeout_s: Optional[float] = None):
        import torch
        import torch.distributed as dist

        # How long a control-plane collective waits for a rank that has really died (not the simulated failures, which
        # keep answering): the reference waits for ever in MPI recv (data_parallelism_train.py:118,135); here the
        # survivors get a PeerLost error after DNNB200_COMM_TIMEOUT_S (default 600 s) instead of hanging.
        if timeout_s is None:
            timeout_s = float(os.environ.get("DNNB200_COMM_TIMEOUT_S", "600"))
        self.timeout_s = timeout_s
        self._dist = dist
        self.rank, self.world, self.local_rank = rank, world, local_rank
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
        os.environ["RANK"], os.environ["WORLD_SIZE"] = str(rank), str(world)
        self._owns = not dist.is_initialized()
        if self._owns:
            backend = "cpu:gloo,cuda:nccl" if torch.cuda.is_available() else "gloo"
            dist.init_process_group(backend=backend, init_method="env://", rank=rank, world_size=world,
                                    timeout=datetime.timedelta(seconds=timeout_s))
        # object traffic always goes over gloo: it must work with ranks that have left the GPU group
        self._cpu_group = (dist.new_group(backend="gloo", timeout=datetime.timedelta(seconds=timeout_s))
                           if dist.get_backend() != "gloo" else dist.group.WORLD)

    def _guard(self, what: str, fn):
        try:
            return fn()
        except RuntimeError as e:        # DistBackendError and gloo's own errors are RuntimeErrors
            raise PeerLost(f"control plane: {what} on rank {self.rank} of {self.world} failed or timed out after "
                           f"{self.timeout_s:.0f} s -- a rank has died or hangs ({str(e).splitlines()[0][:200]})") from e

    def barrier(self):
        self._guard("barrier", lambda: self._dist.barrier(group=self._cpu_group))

    def allgather(self, obj):
        out: List[Any] = [None] * self.world
        self._guard("all-gather", lambda: self._dist.all_gather_object(out, obj, group=self._cpu_group))
        return out

    def bcast(self, obj, root=0):
        box = [obj]
        self._guard("broadcast", lambda: self._dist.broadcast_object_list(box, src=root, group=self._cpu_group))
        return box[0]

    def allreduce_sum_(self, t):
        if t.is_cuda:
            self._guard("all-reduce", lambda: self._dist.all_reduce(t))
        else:
            self._guard("all-reduce", lambda: self._dist.all_reduce(t, group=self._cpu_group))
        return t

    def close(self):
        if self._owns and self._dist.is_initialized():
            self._dist.destroy_process_group()


def init(prefer_mpi: bool = True) -> Comm:
    """The communicator for this process (mpi4py if importable and running under MPI, else env/torch)."""
    rank, world, local = discover()
    if prefer_mpi and any(os.environ.get(v) for v in ("OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "PMIX_RANK")):
        try:
            return _MpiComm()
        except ImportError:
            pass  # mpiexec launched us but mpi4py is absent: the env variables are enough
    if world == 1:
        return Comm()
    return _TorchComm(rank, world, local)
