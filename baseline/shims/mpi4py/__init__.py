"""Minimal stand-in for the slice of mpi4py that is used here -- the image has no MPI runtime.

The reference only calls ``MPI.COMM_WORLD``, ``Get_rank``, ``Get_size`` and the lowercase
pickling ``comm.send(obj, dest)`` / ``comm.recv()``. This shim provides that over
local TCP sockets so the UNMODIFIED reference script can run for the benchmark's reference arm.
It also provides the four collectives the framework's own mpi4py control plane uses
(``Barrier``, ``allgather``, ``bcast``, ``Allreduce`` -- parallel/bootstrap.py ``_MpiComm``), built on
the same point-to-point messages with reserved tags, so that provider can be exercised by a test on a
box without MPI (tests/test_trainer_cpu.py). It is not part of the product: gradients travel
through peer memory, never through this.

Rank/size come from REFSHIM_RANK / REFSHIM_SIZE (else OMPI_COMM_WORLD_RANK / _SIZE, else RANK /
WORLD_SIZE), ports from REFSHIM_PORT + rank. Every user-level send and recv is timestamped so the
harness can time the reference's epochs without touching its code.
"""
import os
import pickle
import queue
import socket
import struct
import threading
import time

_EVENTS = []  # (what, time, nbytes)


def _env_int(names, default):
    for n in names:
        v = os.environ.get(n)
        if v not in (None, ""):
            return int(v)
    return default


_TAG_COLL = -7701   # reserved for the collectives below; user traffic (tag >= 0) never matches it


class _Comm:
    def __init__(self):
        self.rank = _env_int(("REFSHIM_RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "RANK"), 0)
        self.size = _env_int(("REFSHIM_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "WORLD_SIZE"), 1)
        self.base = int(os.environ.get("REFSHIM_PORT", "29700"))
        self._q = queue.Queue()
        self._held = []        # messages taken off the queue that did not match the (source, tag) asked for
        self._srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._srv.bind(("127.0.0.1", self.base + self.rank))
        self._srv.listen(64)
        threading.Thread(target=self._accept_loop, daemon=True).start()

    def _accept_loop(self):
        while True:
            try:
                c, _ = self._srv.accept()
            except OSError:
                return
            threading.Thread(target=self._read_one, args=(c,), daemon=True).start()

    def _read_one(self, c):
        with c:
            hdr = self._read_n(c, 8)
            (n,) = struct.unpack("<Q", hdr)
            self._q.put(self._read_n(c, n))

    @staticmethod
    def _read_n(c, n):
        buf = bytearray()
        while len(buf) < n:
            chunk = c.recv(min(1 << 20, n - len(buf)))
            if not chunk:
                raise ConnectionError("peer closed")
            buf += chunk
        return bytes(buf)

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.size

    def send(self, obj, dest, tag=0):
        payload = pickle.dumps((self.rank, tag, obj), protocol=pickle.HIGHEST_PROTOCOL)
        if tag != _TAG_COLL:
            _EVENTS.append(("send", time.time(), len(payload)))
        deadline = time.time() + 120
        while True:
            try:
                s = socket.create_connection(("127.0.0.1", self.base + dest), timeout=30)
                break
            except OSError:
                if time.time() > deadline:
                    raise
                time.sleep(0.05)
        with s:
            s.sendall(struct.pack("<Q", len(payload)) + payload)

    def recv(self, source=None, tag=None):
        """source/tag None = ANY, as the reference uses it (data_parallelism_train.py:135,227) -- except that ANY
        never matches the collectives' reserved tag."""
        def matches(src, tg):
            if tag is None:
                ok_tag = tg != _TAG_COLL
            else:
                ok_tag = tg == tag
            return ok_tag and (source is None or src == source)

        for k, (src, tg, obj, n) in enumerate(self._held):
            if matches(src, tg):
                self._held.pop(k)
                break
        else:
            while True:
                payload = self._q.get()
                src, tg, obj = pickle.loads(payload)
                n = len(payload)
                if matches(src, tg):
                    break
                self._held.append((src, tg, obj, n))
        if tg != _TAG_COLL:
            _EVENTS.append(("recv", time.time(), n))
        return obj

    # ---- collectives (what parallel/bootstrap.py's _MpiComm calls), rooted at rank 0 ----------------
    def allgather(self, obj):
        if self.size == 1:
            return [obj]
        if self.rank == 0:
            out = [obj] + [None] * (self.size - 1)
            for src in range(1, self.size):
                out[src] = self.recv(source=src, tag=_TAG_COLL)
            for dst in range(1, self.size):
                self.send(out, dst, tag=_TAG_COLL)
            return out
        self.send(obj, 0, tag=_TAG_COLL)
        return self.recv(source=0, tag=_TAG_COLL)

    def Barrier(self):
        self.allgather(None)

    def bcast(self, obj, root=0):
        if self.size == 1:
            return obj
        if self.rank == root:
            for dst in range(self.size):
                if dst != root:
                    self.send(obj, dst, tag=_TAG_COLL)
            return obj
        return self.recv(source=root, tag=_TAG_COLL)

    def Allreduce(self, sendbuf, recvbuf, op=None):
        """Buffer form for numpy arrays; only SUM (the one operation the control plane needs)."""
        import numpy as np

        if op not in (None, SUM):
            raise NotImplementedError("mpi4py shim: Allreduce supports MPI.SUM only")
        parts = self.allgather(np.asarray(sendbuf))
        total = parts[0].copy()
        for p in parts[1:]:          # rank order on every rank: identical bits everywhere
            total = total + p
        np.copyto(recvbuf, total)


SUM = "SUM"


class _MPI:
    _comm = None
    SUM = SUM

    @property
    def COMM_WORLD(self):
        if _MPI._comm is None:
            _MPI._comm = _Comm()
        return _MPI._comm


MPI = _MPI()
