"""data_parallel / heuristic / firing rule (reference: datamodule.lua:68-78,102,235-267)."""
import torch
from hypothesis import given, settings, strategies as st

import tadnn
from tadnn import datamodule, mpi
from oracle import reference_oracle as O


class _H:
    def __init__(self, r, p):
        self._r, self._p = r, p

    def rank(self):
        return self._r

    def size(self):
        return self._p


@settings(max_examples=200, deadline=None)
@given(size=st.integers(1, 5000), world=st.integers(1, 8))
def test_shards_disjoint_equal_and_cover_prefix(size, world):
    data = torch.arange(size)
    seen = []
    lens = set()
    for r in range(world):
        shard, n = datamodule.data_parallel(data, size, _H(r, world), verbose=False)
        assert n == size // world == shard.shape[0]
        lens.add(n)
        seen += shard.tolist()
    assert len(lens) == 1
    assert seen == list(range(world * (size // world)))  # size mod P trailing samples dropped (Q6)


@settings(max_examples=100, deadline=None)
@given(size=st.integers(1, 3000), world=st.integers(1, 8))
def test_matches_oracle_index_arithmetic(size, world):
    data = torch.arange(size * 2).reshape(size, 2)
    for r in range(world):
        a, na = datamodule.data_parallel(data, size, _H(r, world), verbose=False)
        b, nb = O.data_parallel(data, size, _H(r, world))
        assert na == nb and torch.equal(a, b)


def test_shard_is_a_view_not_a_copy():
    data = torch.zeros(10, 3)
    shard, _ = datamodule.data_parallel(data, 10, _H(1, 2), verbose=False)
    shard.fill_(7)
    assert data[5:].eq(7).all() and data[:5].eq(0).all()


def test_remainder_last_and_shuffle():
    data = torch.arange(11)
    s, n = datamodule.data_parallel(data, 11, _H(2, 3), remainder="last", verbose=False)
    assert s.tolist() == [6, 7, 8, 9, 10] and n == 5
    s, n = datamodule.data_parallel(data, 11, _H(1, 3), shuffle=True, verbose=False)
    assert s.tolist() == [1, 4, 7] and n == 3


def test_log_line_matches_reference_format(capsys):
    datamodule.data_parallel(torch.arange(10), 10, _H(1, 2))
    assert capsys.readouterr().out.strip() == "Rank: 1 Start Point: 6 End Point: 10 Stripe: 5 Remainder: 0"


def test_heuristic_table():
    for size, want in [(1, 1), (999, 1), (1000, 10), (2499, 10), (2500, 50), (4999, 50), (5000, 100), (10 ** 6, 100)]:
        assert datamodule.heuristic_batch_size(size) == want == O.heuristic(size)


@settings(max_examples=200, deadline=None)
@given(counter=st.integers(1, 10000), bs=st.integers(1, 200), size=st.integers(1, 10000))
def test_firing_rule(counter, bs, size):
    from tadnn.parallel.dp import SyncState
    s = SyncState.__new__(SyncState)
    s.batchSize, s.size = bs, size
    assert s.should_sync(counter) == O.fires(counter, bs, size)


def test_parallelize_usage_and_minus_one(capsys):
    assert tadnn.parallelize(None, None, None, None) == -1
    assert "-usage for parallelize(data, targets, model, size)" in capsys.readouterr().out


def test_select_collective_lookup():
    w = mpi.FakeWorld(1)
    h = w.handle(0)
    f = datamodule.selectCollective(torch.zeros(3), "sync", "allreduceTensor", h)
    assert callable(f)
    try:
        datamodule.selectCollective(torch.zeros(3), "sync", "nope", h)
        assert False
    except AssertionError as e:
        assert "Could not find collective cpu singlenode sync nope" in str(e)
