"""Preprocessing (cifar example :182-216) and checkpoint/resume utilities."""
import os

import torch

import tadnn
from tadnn import mpi, nn
from tadnn.utils import checkpoint, preprocess


def test_rgb2yuv_and_contrast_norm():
    x = torch.rand(8, 3, 32, 32)
    yuv = preprocess.rgb2yuv(x)
    assert torch.allclose(yuv[:, 0], 0.299 * x[:, 0] + 0.587 * x[:, 1] + 0.114 * x[:, 2], atol=1e-6)
    k = preprocess.gaussian1d(7)
    assert k.numel() == 7 and torch.allclose(k, k.flip(0)) and float(k.max()) == float(k[3])
    tr, te = preprocess.preprocess(x, x[:2])
    assert tr.shape == x.shape and te.shape == (2, 3, 32, 32)
    assert abs(float(tr[:, 1].mean())) < 1e-3 and abs(float(tr[:, 1].std()) - 1) < 1e-2
    assert float(tr[:, 0].mean().abs()) < 0.2  # locally mean-subtracted luminance
    flat = torch.full((1, 1, 32, 32), 0.7)
    assert float(preprocess.spatial_contrastive_normalization(flat, k).abs().max()) < 1e-2  # fp32 round-off / threshold
    trn, _ = preprocess.preprocess(x, None, reference_sign=True)  # the reference divides by -std (Q10)
    assert torch.allclose(trn[:, 1:], -tr[:, 1:], atol=1e-5)


def test_checkpoint_roundtrip_keeps_flat_views_and_counters(tmp_path):
    world = mpi.FakeWorld(2)
    data, tgt = torch.randn(8, 1, 4, 4), torch.randn(8, 1, 3)

    def body(h):
        m = nn.Sequential().add(nn.Reshape(16)).add(nn.Linear(16, 3))
        d, t, n = tadnn.parallelize(data, tgt, m, 8, h, None, 2, verbose=False)
        tr = nn.StochasticGradient(m, nn.MSECriterion())
        tr.learningRate, tr.maxIteration, tr.shuffleIndices, tr.verbose = 0.05, 1, False, False
        tr.train((d, t))
        p = str(tmp_path / "ck.pt")
        checkpoint.save(p, m, tr, mpi=h)
        before = [q.detach().clone() for q in m.parameters()]
        m2 = nn.Sequential().add(nn.Reshape(16)).add(nn.Linear(16, 3))
        tr2 = nn.StochasticGradient(m2, nn.MSECriterion())
        checkpoint.load(p, m2, tr2, mpi=h)
        same = all(torch.equal(a, b) for a, b in zip(before, m2.parameters()))
        return same, tr2.sync_counter, tr.sync_counter

    for same, c2, c1 in world.run(body):
        assert same and c2 == c1 == 5


def test_checkpoint_per_rank_for_model_parallel(tmp_path):
    world = mpi.FakeWorld(2)

    def body(h):
        torch.manual_seed(h.rank())
        m = nn.Sequential().add(nn.MPInitialLinear(8, 4)).add(nn.MPTanh()).add(nn.MPBaseLinear(4, 2))
        p = str(tmp_path / "mp.pt")
        out = checkpoint.save(p, m, mpi=h, per_rank=True)
        assert out.endswith(f".rank{h.rank()}")
        m2 = nn.Sequential().add(nn.MPInitialLinear(8, 4)).add(nn.MPTanh()).add(nn.MPBaseLinear(4, 2))
        checkpoint.load(p, m2, mpi=h, per_rank=True)
        x = torch.randn(3, 8, generator=torch.Generator().manual_seed(0))
        return (m(x) - m2(x)).abs().max().item()

    assert max(world.run(body)) == 0.0


def test_cifar10_t7_batches_roundtrip_and_errors(tmp_path):
    """The on-disk format of the reference's dataset (sgd-torchad_nn-cifar.lua:133-169): ASCII .t7 tables,
    data [3072, n] bytes with one image per column, labels [1, n]."""
    import pytest
    from tadnn.utils import cifar, t7
    g = torch.Generator().manual_seed(0)
    trx = torch.randint(0, 256, (250, 3, 32, 32), generator=g, dtype=torch.uint8)
    try_ = torch.randint(0, 10, (250,), generator=g)
    tex = torch.randint(0, 256, (60, 3, 32, 32), generator=g, dtype=torch.uint8)
    tey = torch.randint(0, 10, (60,), generator=g)
    d = cifar.write_cifar10_t7(str(tmp_path), trx, try_, tex, tey, batch=100)
    raw = t7.load(os.path.join(d, "data_batch_2.t7"), "ascii")
    assert tuple(raw["data"].shape) == (3072, 100) and raw["data"].dtype == torch.uint8  # one image per column
    assert torch.equal(raw["data"][:, 0], trx[100].reshape(-1))
    a, b, c, e = cifar.load_cifar10_t7(str(tmp_path), 230, 50)  # spans three training batches
    assert torch.equal(a, trx[:230].float()) and torch.equal(b, try_[:230])
    assert torch.equal(c, tex[:50].float()) and torch.equal(e, tey[:50]) and e.dtype == torch.int64
    with pytest.raises(ValueError):
        cifar.load_cifar10_t7(str(tmp_path), 251, 50)
    with pytest.raises(FileNotFoundError):
        cifar.load_cifar10_t7(str(tmp_path / "nowhere"), 10, 10)
