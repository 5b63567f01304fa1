"""Timers, the reference's 3-line timing files, a per-epoch JSONL, and the Neptune sink.

Parity:
  * five cumulative wall-clock accumulators (reference: data_parallelism_train.py:32-37)
  * parent file ``log/bs{bs}_log_epochs{E}_proc{P}_parent.txt`` with the three labels of
    reference :126-129, children file ``..._children.txt`` with those of :149-152 -- same file
    names, same label text, one float per line, written once at the end of the run
  * Neptune series ``parameters``, ``train/loss``, ``val/loss``, ``val/acc`` (reference
    :106-112,180-181,250) through an OPTIONAL sink: the package is absent on the box and
    no credential lives in this repository (the reference's hard-coded token is not copied)

New: every epoch also appends one JSON line with device-timed step/exchange figures.
"""
from __future__ import annotations

import json
import os
import time
from contextlib import contextmanager
from dataclasses import dataclass
from typing import Any, Dict, Optional


@dataclass
class Timers:
    data_loading_time: float = 0.0
    training_time: float = 0.0
    evaluation_time: float = 0.0
    mpi_communication_time_parent: float = 0.0
    mpi_communication_time_children: float = 0.0

    @contextmanager
    def add(self, name: str, sync=None):
        """Accumulate wall time into ``name``; ``sync`` (e.g. torch.cuda.synchronize) brackets
        the region so queued GPU work is charged to the right bucket."""
        if sync is not None:
            sync()
        t0 = time.time()
        try:
            yield
        finally:
            if sync is not None:
                sync()
            setattr(self, name, getattr(self, name) + (time.time() - t0))


def log_filename(bs, epochs, nb_proc, role: str) -> str:
    """``bs{bs}_log_epochs{epochs}_proc{nb_proc}_{parent|children}.txt`` (reference :103,144)."""
    assert role in ("parent", "children")
    return f"bs{bs}_log_epochs{epochs}_proc{nb_proc}_{role}.txt"


def write_parent_log(t: Timers, bs, epochs, nb_proc, log_dir: str = "log") -> str:
    os.makedirs(log_dir, exist_ok=True)  # the reference assumes the directory exists
    path = os.path.join(log_dir, log_filename(bs, epochs, nb_proc, "parent"))
    with open(path, "w") as f:
        f.write("Eval data loading time: {0}\n".format(t.data_loading_time))
        f.write("Time spent on evaluation: {0}\n".format(t.evaluation_time))
        f.write("Time spent on parent communication and param sync: {0}\n".format(t.mpi_communication_time_parent))
    return path


def write_children_log(t: Timers, bs, epochs, nb_proc, log_dir: str = "log") -> str:
    os.makedirs(log_dir, exist_ok=True)
    path = os.path.join(log_dir, log_filename(bs, epochs, nb_proc, "children"))
    with open(path, "w") as f:
        f.write("Train data loading time: {0}\n".format(t.data_loading_time))
        f.write("Time spent on training: {0}\n".format(t.training_time))
        f.write("Time spent on children communication: {0}\n".format(t.mpi_communication_time_children))
    return path


def print_parent_summary(t: Timers) -> None:
    print("Eval data loading time: {0}".format(t.data_loading_time))
    print("Time spent on evaluation: {0}".format(t.evaluation_time))
    print("Time spent on parent communication and param sync: {0}".format(t.mpi_communication_time_parent))


def print_children_summary(t: Timers) -> None:
    print("Training data loading time: {0}".format(t.data_loading_time))
    print("Time spent on training: {0}".format(t.training_time))
    print("Time spent on children communication: {0}".format(t.mpi_communication_time_children))


class JsonlLog:
    """One JSON object per line; used for the per-epoch device-timed metrics."""

    def __init__(self, path: Optional[str]):
        self.path = path
        if path:
            os.makedirs(os.path.dirname(path) or ".", exist_ok=True)

    def write(self, **rec: Any) -> None:
        if not self.path:
            return
        rec.setdefault("ts", time.time())
        with open(self.path, "a") as f:
            f.write(json.dumps(rec) + "\n")


# ---------------------------------------------------------------------------
# Neptune: optional, credential-free
# ---------------------------------------------------------------------------
class _Series:
    def __init__(self, store: list):
        self._store = store

    def append(self, v) -> None:
        self._store.append(float(v))

    log = append  # older neptune spelling


class NullRun:
    """Stand-in for ``neptune.init_run(...)``: same ``run[key] = v`` / ``run[key].append(v)`` /
    ``run.stop()`` surface, keeps the values in memory so callers and tests can read them."""

    def __init__(self):
        self.values: Dict[str, Any] = {}
        self.series: Dict[str, list] = {}

    def __setitem__(self, key: str, value) -> None:
        self.values[key] = value

    def __getitem__(self, key: str) -> _Series:
        return _Series(self.series.setdefault(key, []))

    def stop(self) -> None:
        pass


def init_run(project: Optional[str] = None, enabled: bool = False):
    """A Neptune run when asked for AND possible (package importable, ``NEPTUNE_API_TOKEN`` in
    the environment); otherwise a NullRun. Never reads a token from source."""
    if enabled and os.environ.get("NEPTUNE_API_TOKEN"):
        try:
            import neptune  # type: ignore

            return neptune.init_run(project=project or os.environ.get("NEPTUNE_PROJECT"))
        except Exception as e:  # offline box, package missing, ...
            print(f"[neptune] disabled: {e}")
    return NullRun()
