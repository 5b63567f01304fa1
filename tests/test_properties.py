"""Property tests (hypothesis) of the host-side contracts: sharding, batching, the flat layout and membership.

Reference behaviour being pinned: data_parallelism_train.py:49-53 (partition_dataset), DataLoader semantics of
the training loops (:193, drop_last=False), and the 10-key state_dict of models/model.py:11-18.
"""
import torch
from hypothesis import given, settings
from hypothesis import strategies as st

from distributed_neural_network_b200.ops import flat as L
from distributed_neural_network_b200.parallel.membership import Membership
from distributed_neural_network_b200.utils import data as D


class _Range:
    """A dataset that is just its indices (partition functions only need len())."""

    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return i


@settings(max_examples=200, deadline=None)
@given(n=st.integers(1, 60000), size=st.integers(2, 33))
def test_partition_dataset_is_the_reference_formula(n, size):
    ds = _Range(n)
    per = n // (size - 1)
    seen = []
    for rank in range(1, size):
        sub = D.partition_dataset(ds, rank, size)
        assert len(sub) == per
        assert list(sub.indices) == list(range((rank - 1) * per, rank * per))
        seen += list(sub.indices)
    # children shards are disjoint, contiguous, in rank order, and drop only the remainder
    assert seen == list(range(per * (size - 1)))
    assert n - len(seen) < size - 1


@settings(max_examples=100, deadline=None)
@given(n=st.integers(1, 5000), world=st.integers(1, 16))
def test_partition_symmetric_covers_equal_shards(n, world):
    ds = _Range(n)
    per = n // world
    seen = []
    for rank in range(world):
        sub = D.partition_symmetric(ds, rank, world)
        assert len(sub) == per
        seen += list(sub.indices)
    assert seen == list(range(per * world))


def _tiny_dataset(n):
    x = torch.arange(n, dtype=torch.float32).view(n, 1, 1, 1).expand(n, 3, 32, 32).contiguous()
    return D.TensorDataset(x, torch.arange(n) % 10)


@settings(max_examples=40, deadline=None)
@given(n=st.integers(1, 300), bs=st.integers(1, 64), shuffle=st.booleans(), seed=st.integers(0, 2 ** 20))
def test_epoch_tensor_and_batches_visit_every_sample_exactly_once(n, bs, shuffle, seed):
    ds = _tiny_dataset(n)
    g = torch.Generator().manual_seed(seed)
    x, y, tail = D.epoch_tensor(ds, bs, shuffle, g)
    assert x.shape == (n // bs, bs, 3, 32, 32) and y.shape == (n // bs, bs)
    ids = x[:, :, 0, 0, 0].reshape(-1).long().tolist()
    if tail is not None:
        assert 0 < tail[0].shape[0] < bs
        ids += tail[0][:, 0, 0, 0].long().tolist()
    else:
        assert n % bs == 0
    assert sorted(ids) == list(range(n))
    if not shuffle:
        assert ids == list(range(n))
    # the generator-driven iterator agrees with the tensor form (same permutation from the same seed)
    g2 = torch.Generator().manual_seed(seed)
    ids2 = []
    for xb, yb in D.batches(ds, bs, shuffle, g2):
        assert 0 < xb.shape[0] <= bs and torch.equal(yb, xb[:, 0, 0, 0].long() % 10)
        ids2 += xb[:, 0, 0, 0].long().tolist()
    assert ids2 == ids


@settings(max_examples=25, deadline=None)
@given(seed=st.integers(0, 2 ** 20))
def test_flat_layout_roundtrip_and_alignment(seed):
    g = torch.Generator().manual_seed(seed)
    sd = {k: torch.randn(shape, generator=g) for k, shape in L.PARAM_SHAPES.items()}
    flat = L.pack(sd)
    assert flat.numel() == L.FLAT_NUMEL and flat.dtype == torch.float32
    back = L.unpack(flat)
    assert list(back.keys()) == list(L.PARAM_SHAPES.keys())
    for k in sd:
        assert torch.equal(back[k], sd[k])
    mask = L.real_mask()
    assert int(mask.sum()) == sum(v.numel() for v in sd.values()) == 62006
    assert float(flat[~mask].abs().sum()) == 0.0            # alignment gaps stay zero
    for k, off in L.PARAM_OFFSETS.items():
        assert (off * 4) % 128 == 0                          # every tensor starts on a 128-byte boundary
    # views alias the flat buffer
    v = L.views(flat)
    v["fc3.bias"].add_(1.0)
    assert torch.equal(L.unpack(flat)["fc3.bias"], sd["fc3.bias"] + 1.0)


@settings(max_examples=100, deadline=None)
@given(world=st.integers(1, 16), ops=st.lists(st.tuples(st.booleans(), st.integers(0, 15)), max_size=40))
def test_membership_invariants(world, ops):
    m = Membership(world)
    assert m.active == set(range(world)) and m.epoch == 0
    for leave, r in ops:
        r %= world
        before, e0 = set(m.active), m.epoch
        changed = m.leave(r) if leave else m.join(r)
        assert changed == (m.active != before)
        assert m.epoch == e0 + (1 if changed else 0)          # the epoch counts changes of the active set
        assert len(m.active) >= 1                             # the last member never leaves
        assert m.mask == sum(1 << q for q in m.active) and m.divisor == len(m.active)
        assert m.donor() == min(m.active)


def test_donor_is_the_rank_that_kept_the_most_progress():
    """Drop mode: a rank that stayed for the whole epoch donates its state; if every trainer dropped out at some
    point, the one that trained longest does (not whoever has the lowest rank number)."""
    from distributed_neural_network_b200.parallel.membership import Membership

    m = Membership(4)
    assert m.donor() == 0
    # rank 0 dropped at step 3, rank 2 at step 40, ranks 1 and 3 stayed (100 steps)
    assert m.donor({0: (True, 3), 1: (False, 100), 2: (True, 40), 3: (False, 100)}) == 1
    # everybody dropped: rank 2 got furthest
    assert m.donor({0: (True, 0), 1: (True, 17), 2: (True, 40), 3: (True, 5)}) == 2
    # ties go to the lower rank
    assert m.donor({0: (True, 9), 1: (True, 9)}) == 0
