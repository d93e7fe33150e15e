"""Model-parallel layers: forward == dense Linear(cat W_r, sum b_r); exact backward == dense
autograd; compat backward == the literal oracle (reference: nodemodule.lua)."""
import pytest
import torch

from tadnn import mpi, nn
from tadnn.utils import convert
from oracle import reference_oracle as O


def _mp_model(h, compat=False, hidden=32, inp=64, out=10):
    g = torch.Generator().manual_seed(100 + h.rank())
    m = (nn.Sequential().add(nn.MPInitialReshape(inp)).add(nn.MPInitialLinear(inp, hidden, compat=compat))
         .add(nn.MPTanh(compat=compat)).add(nn.MPBaseLinear(hidden, out, compat=compat)))
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(torch.rand(p.shape, generator=g) - 0.5)
    return m


@pytest.mark.parametrize("P", [1, 2, 4])
def test_forward_equals_dense(P):
    world = mpi.FakeWorld(P)
    x = torch.randn(5, 8, 8, generator=torch.Generator().manual_seed(7))

    def body(h):
        m = _mp_model(h)
        dense = convert.dense_model_from_mp(m)
        return m(x).detach(), dense(x).detach(), [c.weight.shape for c in m.children() if hasattr(c, "weight")]

    res = world.run(body)
    for y_mp, y_dense, shapes in res:
        assert torch.allclose(y_mp, y_dense, atol=1e-5)
        assert shapes == [torch.Size([32, 64 // P]), torch.Size([10, 32 // P])]  # [o, i/P] per rank
    for r in range(1, P):
        assert torch.equal(res[0][0], res[r][0])  # replicated output


@pytest.mark.parametrize("P", [2, 4])
def test_exact_backward_equals_dense_autograd(P):
    world = mpi.FakeWorld(P)
    x = torch.randn(6, 8, 8, generator=torch.Generator().manual_seed(8))
    t = torch.randint(0, 10, (6,), generator=torch.Generator().manual_seed(9))

    def body(h):
        m = _mp_model(h)
        dense = convert.dense_model_from_mp(m)
        xi = x.clone().requires_grad_(True)
        loss = torch.nn.functional.cross_entropy(m(xi), t)
        loss.backward()
        xd = x.clone().requires_grad_(True)
        torch.nn.functional.cross_entropy(dense(xd), t).backward()
        r, w1, w2 = h.rank(), 64 // P, 32 // P
        lin = [c for c in dense.children() if isinstance(c, torch.nn.Linear)]
        mpl = [c for c in m.children() if hasattr(c, "weight")]
        errs = [
            (mpl[0].weight.grad - lin[0].weight.grad[:, r * w1:(r + 1) * w1]).abs().max().item(),
            (mpl[1].weight.grad - lin[1].weight.grad[:, r * w2:(r + 1) * w2]).abs().max().item(),
            (mpl[0].bias.grad - lin[0].bias.grad).abs().max().item(),
            (mpl[1].bias.grad - lin[1].bias.grad).abs().max().item(),
            (xi.grad - xd.grad).abs().max().item(),  # full input gradient, gathered
        ]
        return max(errs)

    assert max(world.run(body)) < 1e-5


@pytest.mark.parametrize("P", [2, 4])
def test_compat_backward_equals_oracle_sgd_step(P):
    """One per-sample SGD step through the hooked trainer in compat mode == the literal
    reference chain (Q1 allreduce of gradInput, Q2 tiling, Q3 narrowed MPTanh.output)."""
    world = mpi.FakeWorld(P)
    xs = torch.randn(3, 8, 8, generator=torch.Generator().manual_seed(10))
    ts = torch.randn(3, 10, generator=torch.Generator().manual_seed(11))
    lr = 0.05

    def body(h):
        m = _mp_model(h, compat=True)
        om = O.OSequential(O.OMPInitialReshape(64, mpi=h), O.OMPInitialLinear(64, 32, h), O.OMPTanh(h),
                           O.OMPBaseLinear(32, 10, h))
        lin = [c for c in m.children() if hasattr(c, "weight")]
        om.modules[1].weight, om.modules[1].bias = lin[0].weight.detach().clone(), lin[0].bias.detach().clone()
        om.modules[3].weight, om.modules[3].bias = lin[1].weight.detach().clone(), lin[1].bias.detach().clone()
        crit = nn.MSECriterion()
        ocrit = O.OMSECriterion()
        tr = nn.StochasticGradient(m, crit)
        tr.learningRate, tr.maxIteration, tr.shuffleIndices, tr.verbose = lr, 1, False, False
        tr.train((xs.reshape(3, 1, 8, 8), ts.reshape(3, 1, 10)))
        for i in range(3):
            x, y = xs[i].reshape(1, 8, 8), ts[i].reshape(1, 10)
            ocrit.forward(om.forward(x), y)
            om.updateGradInput(x, ocrit.updateGradInput(om.output, y))
            om.accUpdateGradParameters(x, ocrit.gradInput, lr)
        return max((lin[0].weight.detach() - om.modules[1].weight).abs().max().item(),
                   (lin[0].bias.detach() - om.modules[1].bias).abs().max().item(),
                   (lin[1].weight.detach() - om.modules[3].weight).abs().max().item(),
                   (lin[1].bias.detach() - om.modules[3].bias).abs().max().item())

    assert max(world.run(body)) < 1e-5


def test_compat_differs_from_exact_documenting_q1_q2():
    world = mpi.FakeWorld(2)
    x = torch.randn(4, 8, 8, generator=torch.Generator().manual_seed(12))

    def body(h):
        a, b = _mp_model(h, compat=False), _mp_model(h, compat=True)
        for m in (a, b):
            m(x).square().sum().backward()
        la = [c for c in a.children() if hasattr(c, "weight")][0]
        lb = [c for c in b.children() if hasattr(c, "weight")][0]
        return (la.weight.grad - lb.weight.grad).abs().max().item()

    assert min(world.run(body)) > 1e-4  # the reference's backward is not the true gradient


def test_ctor_asserts_divisibility_and_init_range():
    world = mpi.FakeWorld(4)

    def body(h):
        with pytest.raises(AssertionError):
            nn.MPInitialLinear(10, 3)
        l = nn.MPBaseLinear(64, 7)
        return l.weight.abs().max().item(), l.weight.shape

    for mx, shape in world.run(body):
        assert shape == (7, 16) and mx <= 1 / 4 + 1e-6  # U(+-1/sqrt(i/P)) (Q5)


def test_dense_roundtrip_bias_convention():
    W, b = torch.randn(6, 8), torch.randn(6)
    shards = convert.shard_dense_linear(W, b, 4)
    assert all(s[0].shape == (6, 2) for s in shards) and shards[1][1].abs().sum() == 0
    W2, b2 = convert.merge_mp_linear(shards)
    assert torch.equal(W, W2) and torch.equal(b, b2)


def test_print_dims_switch(monkeypatch, capsys):
    monkeypatch.setenv("TADNN_PRINT_DIMS", "1")
    world = mpi.FakeWorld(1)
    world.run(lambda h: _mp_model(h)(torch.randn(2, 8, 8)))
    out = capsys.readouterr().out
    assert "UPDATE OUTPUT" in out


def test_model_parallelize_rewrites_a_dense_model():
    """tadnn.model_parallelize: the README's by-hand rewrite (README.md:77-110) done automatically; the MP
    model computes what the dense one does and gathers back to the same weights."""
    import tadnn
    from tadnn.utils.convert import dense_model_from_mp
    g = torch.Generator().manual_seed(11)
    dense = nn.Sequential().add(nn.Reshape(16)).add(nn.Linear(16, 12)).add(nn.Tanh()).add(nn.Linear(12, 4))
    with torch.no_grad():
        for p in dense.parameters():
            p.copy_(torch.rand(p.shape, generator=g) - 0.5)
    x = torch.randn(5, 1, 4, 4, generator=g)
    want = dense(x).detach()

    def body(h):
        m = tadnn.model_parallelize(dense, h)
        kinds = [type(c).__name__ for c in m.children()]
        back = dense_model_from_mp(m, h)
        same = all(torch.allclose(a, b) for a, b in zip(back.parameters(), dense.parameters()))
        return kinds, (m(x) - want).abs().max().item(), same

    for kinds, err, same in mpi.FakeWorld(4).run(body):
        assert kinds == ["MPInitialReshape", "MPInitialLinear", "MPTanh", "MPBaseLinear"]
        assert err < 1e-5 and same
