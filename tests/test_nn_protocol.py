"""Torch7 module protocol on top of autograd (tadnn.nn)."""
import torch

from tadnn import nn
from tadnn.utils.confusion import ConfusionMatrix


def _m():
    torch.manual_seed(0)
    return nn.Sequential().add(nn.Reshape(16)).add(nn.Linear(16, 8)).add(nn.Tanh()).add(nn.Linear(8, 3)).add(nn.LogSoftMax())


def test_backward_accumulates_like_torch7():
    m = _m()
    m.needs_grad_input = True
    x = torch.randn(4, 1, 4, 4)
    t = torch.tensor([0, 1, 2, 1])
    crit = nn.ClassNLLCriterion()
    out = m.forward(x)
    loss = crit.forward(out, t)
    g = crit.backward(out.detach(), t)
    m.zeroGradParameters()
    gi = m.backward(x, g)
    m.backward(x, g)  # second call accumulates (gradWeight += ...)
    ref = _m()
    xr = x.clone().requires_grad_(True)
    l = torch.nn.functional.nll_loss(ref(xr), t)
    l.backward()
    assert abs(float(l) - loss) < 1e-6
    assert torch.allclose(gi, xr.grad, atol=1e-6)
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, 2 * q.grad, atol=1e-6)


def test_acc_update_grad_parameters_is_inplace_sgd():
    m, ref = _m(), _m()
    x = torch.randn(1, 4, 4)
    g = torch.randn(1, 3)
    m.forward(x)
    m.updateGradInput(x, g)
    m.accUpdateGradParameters(x, g, 0.1)
    out = ref(x)
    out.backward(g)
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p, q - 0.1 * q.grad, atol=1e-6)
        assert p.grad is None or float(p.grad.abs().sum()) == 0.0  # gradWeight untouched


def test_get_parameters_flattens_and_aliases():
    m = _m()
    flat, gflat = m.getParameters()
    n = sum(p.numel() for p in m.parameters())
    assert flat.numel() >= n and gflat.numel() == flat.numel()
    flat.zero_()
    assert all(float(p.abs().sum()) == 0 for p in m.parameters())


def test_torch7_layer_layouts_and_init():
    c = nn.SpatialConvolutionMM(3, 64, 5, 5)
    assert c.weight.shape == (64, 75) and c.weight.abs().max() <= 1 / 75 ** 0.5 + 1e-6
    assert c(torch.randn(3, 32, 32)).shape == (64, 28, 28)  # unbatched 3-D input
    assert nn.SpatialMaxPooling(3, 3, 3, 3)(torch.randn(2, 64, 28, 28)).shape == (2, 64, 9, 9)
    assert nn.Reshape(1024)(torch.randn(5, 1, 32, 32)).shape == (5, 1024)
    assert nn.Reshape(1024)(torch.randn(1, 32, 32)).shape == (1, 1024)
    assert nn.View(64)(torch.randn(7, 64, 1, 1)).shape == (7, 64)


def test_confusion_matrix():
    cm = ConfusionMatrix(["a", "b", "c"])
    cm.add(torch.tensor([[0.1, 0.8, 0.1], [0.9, 0.05, 0.05]]), torch.tensor([1, 2]))
    cm.add(torch.tensor([0.0, 0.0, 1.0]), 2)
    assert abs(cm.totalValid - 2 / 3) < 1e-9 and "global correct" in str(cm)
