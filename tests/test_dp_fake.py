"""Data-parallel semantics on a FakeWorld (P ranks = P threads): parallelize(), the hooked
backward / StochasticGradient.train, manual mode, vs the literal oracle
(reference: datamodule.lua:15-224)."""
import pytest
import torch

import tadnn
from tadnn import mpi, nn
from oracle import reference_oracle as O


def _mlp(seed=0):
    g = torch.Generator().manual_seed(seed)
    m = nn.Sequential().add(nn.Reshape(16)).add(nn.Linear(16, 12)).add(nn.Tanh()).add(nn.Linear(12, 4))
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(torch.rand(p.shape, generator=g) - 0.5)
    return m


def _omlp(m):
    lin = [c for c in m.children() if isinstance(c, torch.nn.Linear)]
    om = O.OSequential(O.OReshape(16), O.OLinear(16, 12), O.OTanh(), O.OLinear(12, 4))
    om.modules[1].weight, om.modules[1].bias = lin[0].weight.detach().clone(), lin[0].bias.detach().clone()
    om.modules[3].weight, om.modules[3].bias = lin[1].weight.detach().clone(), lin[1].bias.detach().clone()
    return om


DATA = torch.randn(23, 1, 4, 4, generator=torch.Generator().manual_seed(1))
TGT = torch.randn(23, 1, 4, generator=torch.Generator().manual_seed(2))


@pytest.mark.parametrize("P,bs", [(2, 1), (2, 3), (3, 4)])
def test_hooked_stochastic_gradient_equals_oracle(P, bs):
    world = mpi.FakeWorld(P)

    def body(h):
        m = _mlp(seed=5 + h.rank())  # different init per rank: parallelize must broadcast rank 0's
        d, t, n = tadnn.parallelize(DATA, TGT, m, DATA.shape[0], h, None, bs, verbose=False)
        assert n == 23 // P and d.shape[0] == n
        om = _omlp(m)  # after the broadcast
        tr = nn.StochasticGradient(m, nn.MSECriterion())
        tr.learningRate, tr.maxIteration, tr.shuffleIndices, tr.verbose = 0.05, 2, False, False
        tr.train((d, t))
        O.stochastic_gradient_train(om, O.OMSECriterion(), d, t, 0.05, h, bs, n, max_iteration=2)
        lin = [c for c in m.children() if isinstance(c, torch.nn.Linear)]
        err = max((lin[0].weight.detach() - om.modules[1].weight).abs().max().item(),
                  (lin[1].bias.detach() - om.modules[3].bias).abs().max().item())
        return err, tr.syncs, lin[0].weight.detach().clone()

    res = world.run(body)
    n = 23 // P
    expected_syncs = sum(1 for c in range(1, 2 * n + 1) if O.fires(c, bs, n))
    for err, syncs, _ in res:
        assert err < 1e-5 and syncs == expected_syncs
    if bs == 1:  # synced after every sample -> replicas identical
        for r in range(1, P):
            assert torch.allclose(res[0][2], res[r][2], atol=1e-6)


def test_initial_broadcast_makes_replicas_identical():
    world = mpi.FakeWorld(3)

    def body(h):
        m = _mlp(seed=h.rank())
        tadnn.parallelize(DATA, TGT, m, 23, h, None, -1, verbose=False)
        return torch.cat([p.detach().reshape(-1) for p in m.parameters()])

    res = world.run(body)
    assert torch.equal(res[0], res[1]) and torch.equal(res[0], res[2])


def test_manual_mode_and_synchronize_model_averages_params_and_grads():
    world = mpi.FakeWorld(2)

    def body(h):
        m = _mlp(seed=0)
        d, t, n = tadnn.parallelize(DATA, TGT, m, 23, h, None, -1, verbose=False)
        assert getattr(m, "_tadnn_sync", None) is None  # batchSize=-1: no hooks (datamodule.lua:45)
        with torch.no_grad():
            for p in m.parameters():
                p.add_(float(h.rank()))
                p.grad = torch.full_like(p, float(h.rank() + 1))
        tadnn.synchronizeModel(m, h)
        return [p.detach().clone() for p in m.parameters()], [p.grad.clone() for p in m.parameters()]

    (p0, g0), (p1, g1) = world.run(body)
    base = [p.detach() for p in _mlp(seed=0).parameters()]
    for a, b, c, ga in zip(p0, p1, base, g0):
        assert torch.allclose(a, b) and torch.allclose(a, c + 0.5, atol=1e-6)
        assert torch.allclose(ga, torch.full_like(ga, 1.5))


@pytest.mark.parametrize("style", ["t7_backward", "autograd"])
def test_hooked_backward_counts_calls_and_syncs(style):
    world = mpi.FakeWorld(2)
    bs = 3

    def body(h):
        m = _mlp(seed=0)
        d, t, n = tadnn.parallelize(DATA, TGT, m, 23, h, None, bs, verbose=False)
        crit = nn.MSECriterion()
        before = world.calls["allreduce"]
        for i in range(7):
            x, y = d[i:i + 1], t[i:i + 1]
            m.zeroGradParameters()
            if style == "t7_backward":
                out = m.forward(x)
                crit.forward(out, y)
                m.backward(x, crit.backward(out.detach(), y))
            else:
                crit(m(x), y).backward()
            m.updateParameters(0.01)
        h.barrier()
        return m._tadnn_sync.syncs, m.sync_counter, before

    res = world.run(body)
    for syncs, counter, _ in res:
        assert syncs == 2 and counter == 8  # fired at calls 3 and 6; counter starts at 1


def test_grads_only_mode(monkeypatch):
    monkeypatch.setenv("TADNN_SYNC_PROTOTYPE", "0")
    world = mpi.FakeWorld(2)

    def body(h):
        m = _mlp(seed=0)
        tadnn.parallelize(DATA, TGT, m, 23, h, None, 1, verbose=False)
        assert m._tadnn_sync.sync_mode == "grads"
        with torch.no_grad():
            for p in m.parameters():
                p.add_(float(h.rank()))  # params diverge; grads-only sync must not touch them
        x = DATA[:2]
        m(x).sum().backward()
        return [p.detach().clone() for p in m.parameters()], [p.grad.clone() for p in m.parameters()]

    (p0, g0), (p1, g1) = world.run(body)
    assert not torch.allclose(p0[0], p1[0])
    for a, b in zip(g0, g1):
        assert torch.allclose(a, b)


def test_remove_hooks():
    world = mpi.FakeWorld(2)

    def body(h):
        m = _mlp(seed=0)
        tadnn.parallelize(DATA, TGT, m, 23, h, None, 1, verbose=False)
        tadnn.remove_hooks(m)
        n0 = world.calls["allreduce"]
        m(DATA[:1]).sum().backward()
        h.barrier()
        return world.calls["allreduce"] - n0

    assert world.run(body) == [0, 0]


def test_comm_speed_is_live():
    world = mpi.FakeWorld(2)

    def body(h):
        m = _mlp(seed=0)
        return tadnn.comm_speed(DATA, TGT, m, h, None, criterion=nn.MSECriterion(), iters=3)

    assert all(s > 0 for s in world.run(body))


def test_sync_period_from_speed_table():
    from tadnn.datamodule import sync_period_from_speed
    assert sync_period_from_speed(1e-3, 1e-4, 1000) == 1          # sync is already 10 % of one sample
    assert sync_period_from_speed(1e-3, 1e-3, 1000) == 10         # ten samples amortise it to 10 %
    assert sync_period_from_speed(1e-4, 5e-3, 1000, overhead=0.05) == 1000  # clamped to the shard
    assert sync_period_from_speed(0.0, 1e-3, 3000) == 50          # no measurement: the reference's table


def test_parallelize_auto_period_is_measured_and_agreed():
    """batchSize="auto": the period comes from comm_speed(split=True) (datamodule.lua:40-42 is the dead
    call this revives); every rank must end up with the same period, and the model must be untouched."""
    world = mpi.FakeWorld(3)

    def body(h):
        m = _mlp(seed=7)
        before = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).clone()
        d, t, n = tadnn.parallelize(DATA, TGT, m, DATA.shape[0], h, None, "auto", verbose=False,
                                    criterion=nn.MSECriterion())
        after = torch.cat([p.detach().reshape(-1) for p in m.parameters()])
        st = m._tadnn_sync
        return st.batchSize, n, torch.equal(before, after), all(p.grad is None or float(p.grad.abs().sum()) == 0 for p in m.parameters())

    res = world.run(body)
    periods = {r[0] for r in res}
    assert len(periods) == 1 and 1 <= res[0][0] <= res[0][1]
    assert all(r[2] and r[3] for r in res)

    def body2(h):
        return tadnn.comm_speed(DATA, TGT, _mlp(seed=7), h, None, criterion=nn.MSECriterion(), iters=4, split=True, verbose=False)

    out = mpi.FakeWorld(2).run(body2)
    assert out[0] == out[1] and out[0][0] > 0 and out[0][1] > 0  # MAX over ranks: identical on every rank
