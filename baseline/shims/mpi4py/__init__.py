"""Minimal stand-in for the slice of mpi4py the reference uses -- the image has no MPI runtime.

The reference only calls ``MPI.COMM_WORLD``, ``Get_rank``, ``Get_size`` and the lowercase
pickling ``comm.send(obj, dest)`` / ``comm.recv()``. This shim provides exactly
that over local TCP sockets so the UNMODIFIED reference script can run for the benchmark's
reference arm. It is not part of the product: the framework's own exchange is peer memory.

Rank/size come from REFSHIM_RANK / REFSHIM_SIZE, ports from REFSHIM_PORT + rank. Every send and
recv is timestamped so the harness can time the reference's epochs without touching its code.
"""
import os
import pickle
import queue
import socket
import struct
import threading
import time

_EVENTS = []  # (what, time, nbytes)


class _Comm:
    def __init__(self):
        self.rank = int(os.environ.get("REFSHIM_RANK", "0"))
        self.size = int(os.environ.get("REFSHIM_SIZE", "1"))
        self.base = int(os.environ.get("REFSHIM_PORT", "29700"))
        self._q = queue.Queue()
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
        payload = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
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
        payload = self._q.get()
        _EVENTS.append(("recv", time.time(), len(payload)))
        return pickle.loads(payload)


class _MPI:
    _comm = None

    @property
    def COMM_WORLD(self):
        if _MPI._comm is None:
            _MPI._comm = _Comm()
        return _MPI._comm


MPI = _MPI()
