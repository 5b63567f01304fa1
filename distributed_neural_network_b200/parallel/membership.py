"""Membership epochs and failure injection (host side).

Reference behaviour (data_parallelism_train.py:41-46,117,141): once per epoch each rank draws
``rand() < failure_probability`` and, if it "fails", SLEEPS ``failure_duration`` seconds while
everybody else blocks on it. No rank ever leaves.

Here that is ``--failure-mode sleep`` (parity). ``--failure-mode drop`` is the real thing: the
failing rank announces a leave marker through peer memory (``Engine.leave``), the survivors'
exchange kernels drop it from the sum and divide by the survivors from that step on -- mid-epoch,
with no host round trip -- and at the next epoch boundary the group's membership epoch is bumped:
everybody agrees on who is present, returning ranks copy parameters + momentum from a survivor
and adopt its sequence number (``Engine.rejoin``).
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Dict, Optional, Set, Tuple

import numpy as np


@dataclass
class Membership:
    world: int
    epoch: int = 0                       # bumped on every change of the active set
    active: Set[int] = field(default_factory=set)

    def __post_init__(self):
        if not self.active:
            self.active = set(range(self.world))

    @property
    def mask(self) -> int:
        m = 0
        for r in self.active:
            m |= 1 << r
        return m

    @property
    def divisor(self) -> int:
        return len(self.active)

    def leave(self, rank: int) -> bool:
        if rank in self.active and len(self.active) > 1:
            self.active.discard(rank)
            self.epoch += 1
            return True
        return False

    def join(self, rank: int) -> bool:
        if rank not in self.active:
            self.active.add(rank)
            self.epoch += 1
            return True
        return False

    def donor(self, progress: Optional[Dict[int, Tuple[bool, int]]] = None) -> int:
        """The rank a returning member copies its state from.

        ``progress`` maps rank -> (dropped this epoch, steps it completed this epoch). A rank that stayed for the
        whole epoch wins; if EVERY trainer dropped out at some point, the one that trained longest does (so the
        epoch's progress is kept, not the state of whoever happens to have the lowest rank number)."""
        if not progress:
            return min(self.active)
        cands = [r for r in progress if r in self.active or progress[r][0]]
        stayed = [r for r in cands if not progress[r][0]]
        pool = stayed or cands
        return min(pool, key=lambda r: (-progress[r][1], r))


@dataclass
class FailureEvent:
    rank: int
    duration: float
    at_step: Optional[int] = None  # drop mode: the step of the epoch at which the rank leaves


class FailureSimulator:
    """Per-rank, per-epoch Bernoulli failure, like the reference's ``simulate_failure``.

    The reference's RNG is unseeded ``np.random`` (independent per process); pass ``seed`` for
    reproducible drills.
    """

    def __init__(self, probability: float, duration: float, mode: str = "sleep", seed: Optional[int] = None,
                 rank: int = 0):
        if mode not in ("sleep", "drop"):
            raise ValueError("failure mode must be 'sleep' or 'drop'")
        self.p, self.duration, self.mode = float(probability), float(duration), mode
        self._rng = np.random.RandomState(None if seed is None else seed + 7919 * rank)
        self.rank = rank

    def draw(self, steps_in_epoch: int = 0) -> Optional[FailureEvent]:
        if self._rng.rand() < self.p:
            at = int(self._rng.randint(0, max(1, steps_in_epoch))) if self.mode == "drop" else None
            return FailureEvent(self.rank, self.duration, at)
        return None

    def sleep(self, ev: FailureEvent) -> None:
        """The reference's behaviour and messages (data_parallelism_train.py:44-46)."""
        print(f"Process {ev.rank} failed! Sleeping for {ev.duration} seconds.")
        time.sleep(ev.duration)
        print(f"Process {ev.rank} woke up!")
