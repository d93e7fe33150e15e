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


def _scn_by_hand(x, k1, threshold=1e-4, thresval=None):
    """Torch7 nn.SpatialContrastiveNormalization written out with loops (float64): estimator
    E(x)(i,j) = sum_c sum_ab k2[a,b]/(sum k2 * C) * x[c, i+a-p, j+b-p] with zero padding, coef = E(ones),
    subtractive x - E(x)/coef, divisive x / T(sqrt(E(x^2))/coef)."""
    thresval = threshold if thresval is None else thresval
    x = x.double()
    C, H, W = x.shape
    k2 = torch.outer(k1, k1).double()
    k2 = k2 / (k2.sum() * C)
    p = k1.numel() // 2

    def E(z):
        out = torch.zeros(H, W, dtype=torch.float64)
        for i in range(H):
            for j in range(W):
                s = 0.0
                for a in range(k1.numel()):
                    for b in range(k1.numel()):
                        ii, jj = i + a - p, j + b - p
                        if 0 <= ii < H and 0 <= jj < W:
                            s += float(k2[a, b]) * float(z[:, ii, jj].sum())
                out[i, j] = s
        return out

    coef = E(torch.ones_like(x))
    c = x - E(x) / coef
    sd = torch.sqrt(E(c * c)) / coef
    sd = torch.where(sd > threshold, sd, torch.full_like(sd, thresval))
    return c / sd


def test_spatial_contrastive_normalization_is_torch7_arithmetic():
    from tadnn.utils import preprocess
    g = torch.Generator().manual_seed(3)
    k = preprocess.gaussian1d(5)
    x1 = torch.rand(1, 8, 9, generator=g)
    m = nn.SpatialContrastiveNormalization(1, k)
    assert torch.allclose(m(x1).double(), _scn_by_hand(x1, k), atol=1e-5)          # [C, H, W] input, borders included
    assert torch.allclose(m(x1.unsqueeze(0))[0], m(x1))                             # batch mode
    assert torch.allclose(preprocess.spatial_contrastive_normalization(x1.unsqueeze(0), k)[0], m(x1), atol=1e-6)
    x3 = torch.rand(3, 7, 7, generator=g)                                           # several planes share one local mean / std
    assert torch.allclose(nn.SpatialContrastiveNormalization(3, k)(x3).double(), _scn_by_hand(x3, k), atol=1e-5)
    assert torch.allclose(m(4.0 * x1), m(x1), atol=1e-4)                            # contrast invariance above the threshold
    flat = torch.full((1, 8, 8), 0.3)
    assert m(flat).abs().max() < 1e-2                                               # std <= threshold -> divided by thresval
    big = nn.SpatialDivisiveNormalization(1, k, threshold=10.0, thresval=2.0)       # element-wise Threshold(threshold, thresval)
    assert torch.allclose(big(x1), x1 / 2.0)
    sub = nn.SpatialSubtractiveNormalization(1, k)
    assert torch.allclose(sub(flat), torch.zeros_like(flat), atol=1e-6)             # border-corrected mean of a constant is the constant
    import pytest
    with pytest.raises(ValueError):
        nn.SpatialContrastiveNormalization(1, torch.ones(4))                        # even kernels are refused, as in Torch7
