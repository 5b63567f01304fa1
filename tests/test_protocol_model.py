"""An executable model of the exchange protocols of csrc/exchange.cuh (one-shot, two-shot), run under random interleavings.

compute-sanitizer cannot see a race between GPUs, and it is closed on the development pool anyway;
what CAN be checked on a CPU is the protocol itself. The model keeps exactly the state the kernels keep --
per-rank receive areas [parity][source][slice] of LL lines {value, seq}, per-rank flag rows with leave markers --
and exactly their rules:

  * a rank pushes ALL its lines of a step (to every configured peer) or none, then polls its OWN receive area;
  * a line counts when its sequence number equals this step's; sequence numbers only grow, nothing is reset;
  * parity = seq & 1 double-buffers the receive area;
  * a rank that leaves writes a leave marker L (first sequence number it will NOT take part in) into every peer's
    flag row AFTER the pushes of its last step; a waiter treats it as absent for seq >= L;
  * a rank starts step t+1 only when all its slices of step t are summed (the update kernel is one grid).

A scheduler picks the next rank at random after every single store or load, so every interleaving of remote
stores and local polls the hardware could produce at that granularity is reachable. Checked: every rank that
takes part in a step gets the same sum and the same divisor for every slice, they equal the ground truth, and no
line is ever read with another step's payload (the parity argument).

Reference: this replaces the blocking recv star of data_parallelism_train.py:226-244, where the parent's
`comm.recv()` serialises everything and a sleeping child stalls every rank.
"""
import random

import pytest

S = 3          # slices per rank (the real vector has 122; the protocol is per slice)


def payload(rank, step, s):
    return 1000.0 * step + 10.0 * rank + s


class World:
    early_leave_lookup = False   # round 2: markers are also read at the START of a step (update kernel, before its wait)

    def __init__(self, W, steps, leave_at):
        self.W, self.steps = W, steps
        self.leave_at = leave_at                       # rank -> first step (0-based) it does not take part in
        # recv[p][parity][src][slice] = (value, seq); flags[p][src] = leave marker (0 = present)
        self.recv = [[[[(0.0, 0) for _ in range(S)] for _ in range(W)] for _ in range(2)] for _ in range(W)]
        self.leave = [[0] * W for _ in range(W)]
        self.results = {}                              # (rank, step, slice) -> (sum, n)


def rank_program(w: World, r: int):
    """One rank as a generator: yields after every single remote store / local load."""
    W = w.W
    for step in range(w.steps):
        seq = step + 1
        if w.leave_at.get(r) == step:
            # leave: markers go out after the pushes of the last step this rank took part in (already done)
            for p in range(W):
                if p != r:
                    w.leave[p][r] = seq
                    yield
            return
        parity = seq & 1
        # 0. (round 2) look the leave markers up early -- a racy read: a marker that lands later is found by the waiters
        #    below, once, the slow way. A rank whose marker covers this step is not pushed to and not waited for.
        gone = set()
        if w.early_leave_lookup:
            for q in range(W):
                if q != r:
                    L = w.leave[r][q]
                    yield
                    if L != 0 and seq >= L:
                        gone.add(q)
        # 1. push all lines of this step to every peer (a departed peer keeps its memory mapped: harmless)
        for s in range(S):
            for p in range(W):
                if p == r or p in gone:
                    continue
                w.recv[p][parity][r][s] = (payload(r, step, s), seq)
                yield
        # 2. per slice: poll my own receive area until every peer's line carries seq, or the peer has left
        for s in range(S):
            got, absent = {}, set(gone)
            pending = [q for q in range(W) if q != r and q not in gone]
            while pending:
                for q in list(pending):
                    val, sq = w.recv[r][parity][q][s]
                    yield
                    if sq == seq:
                        assert val == payload(q, step, s), "a line of another step was accepted"
                        got[q] = val
                        pending.remove(q)
                        continue
                    L = w.leave[r][q]
                    yield
                    if L != 0 and seq >= L:
                        # it pushed every line of its last step BEFORE the marker: one more look settles it
                        val, sq = w.recv[r][parity][q][s]
                        yield
                        if sq == seq:
                            got[q] = val
                        else:
                            absent.add(q)
                        pending.remove(q)
            total, n = 0.0, 0
            for q in range(W):                          # rank order, own contribution at its position
                if q == r:
                    total += payload(r, step, s); n += 1
                elif q in got:
                    total += got[q]; n += 1
            w.results[(r, step, s)] = (total, n)
    return


def run(W, steps, leave_at, seed, early_leave_lookup=False):
    rng = random.Random(seed)
    w = World(W, steps, leave_at)
    w.early_leave_lookup = early_leave_lookup
    progs = {r: rank_program(w, r) for r in range(W)}
    budget = 2_000_000
    while progs:
        r = rng.choice(list(progs))
        # bursts of random length: lets one rank run far ahead of another, as a fast GPU would
        for _ in range(rng.choice((1, 1, 2, 5, 40))):
            try:
                next(progs[r])
            except StopIteration:
                del progs[r]
                break
            budget -= 1
            assert budget > 0, "livelock: some rank waits for a line that never comes"
    return w


def check(w: World):
    for step in range(w.steps):
        present = [r for r in range(w.W) if w.leave_at.get(r, 10 ** 9) > step]
        for s in range(S):
            want = (sum(payload(r, step, s) for r in present), len(present))
            for r in present:
                assert w.results[(r, step, s)] == want, (r, step, s, w.results[(r, step, s)], want)
    # nobody produced results for steps it did not take part in
    for (r, step, s) in w.results:
        assert w.leave_at.get(r, 10 ** 9) > step


@pytest.mark.parametrize("W", [2, 3, 5, 8])
def test_static_membership_all_interleavings_agree(W):
    for seed in range(60):
        check(run(W, steps=5, leave_at={}, seed=seed))


@pytest.mark.parametrize("W", [3, 4, 8])
def test_ranks_leaving_mid_run_survivors_agree(W):
    rng = random.Random(W)
    for seed in range(80):
        leavers = rng.sample(range(W), k=rng.randint(1, W - 1))
        leave_at = {r: rng.randint(1, 5) for r in leavers}
        check(run(W, steps=7, leave_at=leave_at, seed=1000 + seed))


@pytest.mark.parametrize("W", [3, 4, 8])
def test_early_leave_lookup_keeps_survivors_in_agreement(W):
    """Round 2: the update kernel reads the leave markers before its dependency wait, so survivors skip a departed rank
    from the first instruction instead of polling for it in every step. The early read races with the marker's
    arrival; whoever misses it falls back to the waiters' check. Everybody must still agree on every step."""
    rng = random.Random(50 + W)
    for seed in range(80):
        leavers = rng.sample(range(W), k=rng.randint(1, W - 1))
        leave_at = {r: rng.randint(1, 5) for r in leavers}
        check(run(W, steps=7, leave_at=leave_at, seed=3000 + seed, early_leave_lookup=True))


def rank_program_two_shot(w: World, r: int):
    """The two-shot variant (static membership): slice s is owned by rank s % W; the others send their value only to
    the owner (source slot = sender), the owner adds in rank order and sends the SUM to everybody through its own
    source slot, which nobody else uses for that slice."""
    W = w.W
    for step in range(w.steps):
        seq = step + 1
        parity = seq & 1
        for s in range(S):                              # hop 1: contributions to the owners
            owner = s % W
            if owner != r:
                w.recv[owner][parity][r][s] = (payload(r, step, s), seq)
                yield
        for s in range(S):
            owner = s % W
            if owner == r:                              # owner: gather, add in rank order, broadcast
                vals = {}
                pending = [q for q in range(W) if q != r]
                while pending:
                    for q in list(pending):
                        val, sq = w.recv[r][parity][q][s]
                        yield
                        if sq == seq:
                            assert val == payload(q, step, s), "a line of another step was accepted"
                            vals[q] = val
                            pending.remove(q)
                total = 0.0
                for q in range(W):
                    total += payload(r, step, s) if q == r else vals[q]
                for p in range(W):
                    if p != r:
                        w.recv[p][parity][r][s] = (total, seq)
                        yield
                w.results[(r, step, s)] = (total, W)
            else:                                       # everybody else: wait for the owner's sum
                while True:
                    val, sq = w.recv[r][parity][owner][s]
                    yield
                    if sq == seq:
                        w.results[(r, step, s)] = (val, W)
                        break


@pytest.mark.parametrize("W", [2, 4, 8])
def test_two_shot_all_interleavings_agree(W):
    for seed in range(60):
        rng = random.Random(seed)
        w = World(W, 5, {})
        progs = {r: rank_program_two_shot(w, r) for r in range(W)}
        budget = 2_000_000
        while progs:
            r = rng.choice(list(progs))
            for _ in range(rng.choice((1, 1, 2, 5, 40))):
                try:
                    next(progs[r])
                except StopIteration:
                    del progs[r]
                    break
                budget -= 1
                assert budget > 0, "livelock"
        check(w)


def test_the_model_catches_a_protocol_without_parity():
    """Sanity of the checker itself: with ONE receive buffer instead of two parities, a fast rank's push of step
    t+1 can land before a slow rank has read step t -- the model must notice (a line with another step's seq
    makes the slow rank wait forever, which the livelock budget reports)."""
    # emulate the broken protocol by aliasing both parities of every receive area to the same storage
    broken = 0
    for seed in range(40):
        rng = random.Random(seed)
        w = World(2, 6, {})
        for p in range(2):
            w.recv[p][1] = w.recv[p][0]
        progs = {r: rank_program(w, r) for r in range(2)}
        budget = 200_000
        try:
            while progs:
                r = rng.choice(list(progs))
                for _ in range(rng.choice((1, 2, 5, 40))):
                    try:
                        next(progs[r])
                    except StopIteration:
                        del progs[r]
                        break
                    budget -= 1
                    if budget <= 0:
                        raise AssertionError("livelock")
            check(w)
        except AssertionError:
            broken += 1
    assert broken > 0


# ---------------------------------------------------------------------------------------------------------
# The step-to-step dependency of chained steps: the forward/backward kernel of step t+1 waits on
# a counter the update kernel's last CTA releases, not on the completion of the update grid.
# ---------------------------------------------------------------------------------------------------------
def _run_chained(steps, n_f, n_u, seed, f_waits=True):
    """F CTAs read ALL parameter slices and write their records; U CTAs read ALL records and write ONE slice.
    Grids: U(t) may start only when F(t) has completed (griddepcontrol.wait in the update kernel); F(t+1) CTAs are
    launched as soon as every U(t) CTA has started (programmatic launch) and gate themselves on the counter."""
    rng = random.Random(seed)
    params = [0] * n_u            # version of each slice = number of updates applied
    records = [-1] * n_f          # records[i] = step whose forward/backward wrote it
    counter = [0]
    done = [0]
    state = {"f_done": [0] * (steps + 1), "u_started": [0] * (steps + 1)}

    def f_cta(t, i):
        while state["u_started"][t - 1] < n_u and t > 0:      # not launched before all U(t-1) CTAs have started
            yield
        if f_waits:
            while counter[0] < t:                               # ld.acquire.gpu spin (thread 0)
                yield
        for k in range(n_u):                                    # TMA of the weights
            assert params[k] == t, f"F({t}) read slice {k} at version {params[k]}"
            yield
        records[i] = t
        yield
        state["f_done"][t] += 1

    def u_cta(t, k):
        state["u_started"][t] += 1
        while state["f_done"][t] < n_f:                         # griddepcontrol.wait: F(t) complete
            yield
        for i in range(n_f):
            assert records[i] == t, f"U({t}) read record {i} of step {records[i]}"
            yield
        params[k] = t + 1
        yield
        done[0] += 1
        if done[0] == n_u:                                      # last CTA out: reset, then release the counter
            done[0] = 0
            counter[0] = t + 1

    progs = []
    for t in range(steps):
        progs += [f_cta(t, i) for i in range(n_f)] + [u_cta(t, k) for k in range(n_u)]
    # the stream launches grids in order, but CTAs of already-launched grids interleave freely: keep a window
    live, nxt, budget = [], 0, 1_000_000
    while live or nxt < len(progs):
        while nxt < len(progs) and len(live) < 3 * (n_f + n_u):
            live.append(progs[nxt]); nxt += 1
        g = rng.choice(live)
        try:
            next(g)
        except StopIteration:
            live.remove(g)
        budget -= 1
        assert budget > 0, "deadlock in the chained-step model"
    assert params == [steps] * n_u and counter[0] == steps


def test_chained_step_counter_dependency_is_sufficient():
    for seed in range(150):
        _run_chained(steps=5, n_f=4, n_u=5, seed=seed)


def test_chained_step_model_fails_without_the_counter_wait():
    failures = 0
    for seed in range(60):
        try:
            _run_chained(steps=4, n_f=4, n_u=5, seed=seed, f_waits=False)
        except AssertionError:
            failures += 1
    assert failures > 0


# ---------------------------------------------------------------------------------------------------------
# One-shot with 128-byte lines (exchange.cuh, variant 3): a line is 31 values and ONE flag word, written by one
# warp-wide store and read by one warp-wide load. What the variant assumes of the hardware is that such a store
# becomes visible in the peer's memory as a unit (tests/test_gpu_multiproc.py::test_ll128_lines_are_never_torn hammers
# exactly that over NVLink). The model states the assumption and shows what breaks without it.
# ---------------------------------------------------------------------------------------------------------
LINE_VALUES = 3        # values per line in the model (31 in the kernel)


def rank_program_lines(w, r, atomic_lines, rng):
    """recv128[p][parity][src][line] = [v0, .., v_{n-1}, flag]; jobs (line, peer) are polled in any order."""
    W = w.W
    for step in range(w.steps):
        seq = step + 1
        parity = seq & 1
        for line in range(S):
            words = [payload(r, step, line) + 0.1 * k for k in range(LINE_VALUES)] + [seq]
            for p in range(W):
                if p == r:
                    continue
                dst = w.recv128[p][parity][r][line]
                if atomic_lines:
                    dst[:] = words
                    yield
                else:                                   # a fabric that delivers the words of a line one by one, any order
                    for k in rng.sample(range(len(words)), len(words)):
                        dst[k] = words[k]
                        yield
        jobs = [(line, q) for line in range(S) for q in range(W) if q != r]
        got = {}
        while jobs:
            for job in list(jobs):
                line, q = job
                snap = list(w.recv128[r][parity][q][line])      # one warp-wide load: a snapshot of the whole line
                yield
                if snap[-1] == seq:
                    got[job] = snap[:-1]
                    jobs.remove(job)
        for line in range(S):
            for k in range(LINE_VALUES):
                total = 0.0
                for q in range(W):                              # rank order, own contribution at its position
                    total += payload(r, step, line) + 0.1 * k if q == r else got[(line, q)][k]
                w.results[(r, step, line, k)] = total


def _run_lines(W, steps, seed, atomic_lines):
    rng = random.Random(seed)
    w = World(W, steps, {})
    w.recv128 = [[[[[0.0] * LINE_VALUES + [0] for _ in range(S)] for _ in range(W)] for _ in range(2)] for _ in range(W)]
    progs = {r: rank_program_lines(w, r, atomic_lines, rng) for r in range(W)}
    budget = 2_000_000
    while progs:
        r = rng.choice(list(progs))
        for _ in range(rng.choice((1, 1, 2, 5, 40))):
            try:
                next(progs[r])
            except StopIteration:
                del progs[r]
                break
            budget -= 1
            assert budget > 0, "livelock"
    for step in range(steps):
        for line in range(S):
            for k in range(LINE_VALUES):
                want = sum(payload(q, step, line) + 0.1 * k for q in range(W))
                for r in range(W):
                    assert abs(w.results[(r, step, line, k)] - want) < 1e-9, (r, step, line, k)


@pytest.mark.parametrize("W", [2, 4, 8])
def test_128_byte_lines_with_one_flag_word_agree_when_lines_arrive_whole(W):
    for seed in range(40):
        _run_lines(W, steps=5, seed=seed, atomic_lines=True)


def test_128_byte_lines_need_lines_that_arrive_whole():
    """Negative control: if the words of a line could arrive separately, a reader that sees this step's flag word may
    still see last-but-one step's values next to it -- the sums go wrong. This is why the variant is only enabled
    after the probe kernel has found no torn line on the actual links."""
    broken = 0
    for seed in range(40):
        try:
            _run_lines(3, steps=6, seed=seed, atomic_lines=False)
        except AssertionError:
            broken += 1
    assert broken > 0
