"""CNN kernels and the fused CifarCNN loss vs plain PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def test_im2col_nchw_and_nhwc():
    from tadnn.ops import cnn as K
    x = torch.randn(3, 3, 32, 32, device="cuda").bfloat16()
    col = K.im2col(x, 5, 5, 80, nchw=True)
    ref = F.unfold(x.float(), 5).transpose(1, 2).reshape(3 * 784, 3, 25).permute(0, 2, 1).reshape(3 * 784, 75)
    assert torch.equal(col[:, :75].float(), ref) and float(col[:, 75:].abs().sum()) == 0
    y = torch.randn(2, 9, 9, 64, device="cuda").bfloat16()
    col2 = K.im2col(y, 5, 5, 1600, nchw=False)
    ref2 = F.unfold(y.float().permute(0, 3, 1, 2), 5).transpose(1, 2).reshape(2 * 25, 64, 25).permute(0, 2, 1).reshape(50, 1600)
    assert torch.equal(col2.float(), ref2)


def test_col2im_is_adjoint_of_im2col():
    from tadnn.ops import cnn as K
    dcol = torch.randn(2 * 25, 1600, device="cuda").bfloat16()
    dx = K.col2im(dcol, 2, 64, 9, 9, 5, 5)
    cols = dcol.float().reshape(2, 25, 25, 64).permute(0, 3, 2, 1).reshape(2, 1600, 25)  # [B, C*25 (c,k), L]
    ref = F.fold(cols, (9, 9), 5).permute(0, 2, 3, 1)
    assert (dx.float() - ref).abs().max().item() < 0.05 * ref.abs().max().item()


def test_maxpool_fwd_bwd():
    from tadnn.ops import cnn as K
    x = torch.relu(torch.randn(4, 28, 28, 64, device="cuda")).bfloat16()
    y, idx = K.maxpool_fwd(x, 3)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(xr, 3, 3)
    assert torch.equal(y.float(), yr.permute(0, 2, 3, 1))
    dy = torch.randn(4, 9, 9, 64, device="cuda").bfloat16()
    dx = K.maxpool_relu_bwd(dy, y, idx, 28, 28, 3)
    # reference: route to argmax where pooled value > 0 (ties: first max, like our scan order)
    (gr,) = torch.autograd.grad(yr, xr, (dy.float() * (y.float() > 0)).permute(0, 3, 1, 2))
    # tie-breaking can differ only where the max is 0 (masked) -> exact match expected
    assert torch.equal(dx.float(), gr.permute(0, 2, 3, 1))


def test_colsum_pack_unpack_sgd():
    from tadnn.ops import cnn as K
    x = torch.randn(10000, 64, device="cuda").bfloat16()
    out = torch.ones(64, device="cuda")
    K.colsum_bf16(x, out)
    assert torch.allclose(out, 1 + x.float().sum(0), atol=2e-2, rtol=1e-4)
    w = torch.randn(64, 75, device="cuda")
    wp = K.conv_weight_pack(w, 3, 5, 5, 80)
    ref = w.view(64, 3, 5, 5).permute(0, 2, 3, 1).reshape(64, 75)
    assert torch.equal(wp[:, :75].float(), ref.bfloat16().float()) and float(wp[:, 75:].abs().sum()) == 0
    g = torch.ones(64, 75, device="cuda")
    K.conv_wgrad_unpack(F.pad(ref, (0, 5)).contiguous(), g, 3, 5, 5, 2.0)
    assert torch.allclose(g, 1 + 2 * w)
    p, gr = torch.randn(1000, device="cuda"), torch.randn(1000, device="cuda")
    p0, g0 = p.clone(), gr.clone()
    K.sgd_step(p, gr, 0.1, 0.01, True)
    assert torch.allclose(p, p0 - 0.1 * (g0 + 0.01 * p0), atol=1e-6) and float(gr.abs().sum()) == 0


def test_fused_conv1_matches_unfused_kernels():
    """conv1_pool_forward/backward (implicit GEMM, pooled epilogue) vs im2col + GEMM + pooling kernels."""
    from tadnn.ops import cnn as K
    torch.manual_seed(0)
    for B in (1, 5, 130):
        x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
        w = torch.randn(64, 75, device="cuda") * 0.2
        b = torch.randn(64, device="cuda") * 0.1
        p1, idx1 = K.conv1_pool_forward(x, K.conv1_weight_pack(w), b)
        col = K.im2col(x, 5, 5, 80, nchw=True)
        a1 = K.gemm_bf16(col, K.conv_weight_pack(w, 3, 5, 5, 80), bias=b, act="relu", tile_n=64)
        p1_ref, idx_ref = K.maxpool_fwd(a1.view(B, 28, 28, 64), 3)
        assert (p1.float() - p1_ref.float()).abs().max().item() <= 0.02 * max(1.0, p1_ref.float().abs().max().item())
        same = (idx1 == idx_ref) | (p1_ref <= 0)
        assert same.float().mean().item() > 0.995
        dp1 = torch.randn(B, 9, 9, 64, device="cuda").bfloat16()
        gw = torch.zeros(64, 75, device="cuda")
        K.conv1_pool_backward(x, dp1, p1_ref, idx_ref, gw)
        da1 = K.maxpool_relu_bwd(dp1, p1_ref, idx_ref, 28, 28, 3).view(B * 784, 64)
        dwp = K.gemm_bf16(da1, col, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32, tile_n=128)
        gw_ref = torch.zeros(64, 75, device="cuda")
        K.conv_wgrad_unpack(dwp, gw_ref, 3, 5, 5)
        rel = ((gw - gw_ref).norm() / gw_ref.norm()).item()
        assert rel < 2e-3, (B, rel)


def test_conv2_modes_agree():
    """dense / pooled / legacy conv2 paths give the same loss and the same gradients."""
    from tadnn.models.cifar_cnn import CifarCNN
    res = {}
    for mode in ("legacy", "dense", "pooled"):
        torch.manual_seed(0)
        # conv1_kernel="rows": the three modes then share conv1's rounding points (the sliding-operand conv1 of the
        # default "dense" path breaks bf16 pooling ties by the fp32 value, the unfused conv1 of "legacy" by position)
        m = CifarCNN(conv2_mode=mode, conv1_kernel="rows").cuda().eval()
        x = torch.randn(37, 3, 32, 32, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5)).bfloat16()
        t = torch.arange(37, device="cuda") % 10
        loss = m.loss(x, t)
        loss.backward()
        res[mode] = (loss.item(), {n: p.grad.clone() for n, p in m.named_parameters()})
    for mode in ("dense", "pooled"):
        assert abs(res[mode][0] - res["legacy"][0]) < 1e-6
        for n, g in res[mode][1].items():
            ref = res["legacy"][1][n]
            assert ((g - ref).norm() / (ref.norm() + 1e-12)).item() < 1.5e-2, (mode, n)
    for n, g in res["pooled"][1].items():
        assert torch.equal(g, res["dense"][1][n]) or ((g - res["dense"][1][n]).norm() / (g.norm() + 1e-12)).item() < 1e-3


@pytest.mark.parametrize("fused,mode,kern", [(True, "dense", "sliding"), (True, "dense", "rows"), (True, "pooled", "rows"),
                                             (False, "legacy", "rows")])
@pytest.mark.parametrize("B", [1, 7, 64, 300])
def test_cifar_cnn_loss_and_grads_match_pytorch(B, fused, mode, kern):
    from tadnn.models.cifar_cnn import CifarCNN
    torch.manual_seed(0)
    m = CifarCNN(fused_conv1=fused, conv2_mode=mode, conv1_kernel=kern).cuda().eval()  # eval: dropout off, exact comparison
    with torch.no_grad():  # widen the init a little so ReLUs/pools are well exercised
        for p in m.parameters():
            p.mul_(2.0)
    x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
    t = torch.randint(0, 10, (B,), device="cuda")
    loss, logp = m.loss_and_logp(x, t)
    loss.backward()
    got = {n: p.grad.clone() for n, p in m.named_parameters()}
    m.zero_grad()
    ref_logp = m.reference_logp(x.float(), emulate_bf16=True)
    ref_loss = F.nll_loss(ref_logp, t)
    ref_loss.backward()
    assert abs(loss.item() - ref_loss.item()) < 2e-2 * max(1.0, abs(ref_loss.item()))
    assert (logp - ref_logp.detach()).abs().max().item() < 0.05
    report = {}
    for n, p in m.named_parameters():
        ref = p.grad
        report[n] = ((got[n] - ref).norm() / (ref.norm() + 1e-12)).item()
    print("\n[cnn grads rel-err B=%d] %s" % (B, {k: round(v, 4) for k, v in report.items()}))
    # The error is dominated by arg-max / ReLU decisions that flip on near-ties between the bf16 kernel and the
    # reference (a flipped pooling winner moves a whole gradient contribution); it fluctuates with the random draw
    # (B = 1, 7: < 0.4 %; B = 64, 300: 1 - 5 % for conv1, whose gradient sums the most such decisions).
    for n, v in report.items():
        assert v < (6e-2 if n == "conv1_weight" else 4e-2), f"{n}: relative error {v}"


def test_forward_backward_matches_autograd_path():
    from tadnn.models.cifar_cnn import CifarCNN
    torch.manual_seed(0)
    a, b = CifarCNN().cuda().eval(), CifarCNN().cuda().eval()
    b.load_state_dict(a.state_dict())
    x = torch.randn(50, 3, 32, 32, device="cuda").bfloat16()
    t = torch.randint(0, 10, (50,), device="cuda")
    la = a.loss(x, t)
    la.backward()
    lb = b.forward_backward(x, t)
    lb2 = b.forward_backward(x, t)  # accumulates
    assert la.item() == lb.item() == lb2.item()
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        assert ((q.grad - 2 * p.grad).norm() / (p.grad.norm() + 1e-12)).item() < 2e-3, n


def test_cifar_cnn_dropout_is_consistent_and_seeded():
    from tadnn.models.cifar_cnn import CifarCNN
    torch.manual_seed(0)
    m = CifarCNN().cuda().train()
    x = torch.randn(256, 3, 32, 32, device="cuda").bfloat16()
    t = torch.randint(0, 10, (256,), device="cuda")
    l1 = m.loss(x, t).item()
    l2 = m.loss(x, t).item()
    assert l1 != l2  # a new mask every step
    m._step = 0
    assert m.loss(x, t).item() == l1  # same seed/step -> same mask (deterministic loss reduction)


def test_cifar_cnn_trains():
    from tadnn.models.cifar_cnn import CifarCNN
    from tadnn.ops import cnn as K
    torch.manual_seed(0)
    m = CifarCNN(dropout=0.0).cuda().train()
    t = torch.randint(0, 10, (512,), device="cuda")
    patterns = torch.randn(10, 3, 32, 32, device="cuda")
    x = (0.5 * torch.randn(512, 3, 32, 32, device="cuda") + patterns[t]).bfloat16()  # learnable
    losses = []
    for _ in range(60):
        m.zero_grad()
        loss = m.loss(x, t)
        loss.backward()
        for p in m.parameters():
            K.sgd_step(p.data.view(-1), p.grad.view(-1), 0.05)
        losses.append(loss.item())
    assert min(losses[-10:]) < 0.8 * losses[0], (losses[0], losses[-10:])


def test_per_sample_graph_step_matches_eager_and_is_used_by_stochastic_gradient():
    """The reference's loop is one sample per step (datamodule.lua:134-158): on a GPU each step is one replay of a
    captured CUDA graph of the fused kernels + the in-place update."""
    from tadnn import nn
    from tadnn.models.cifar_cnn import CifarCNN
    from tadnn.ops import cnn as K
    torch.manual_seed(0)
    a, b = CifarCNN(dropout=0.0).cuda().train(), CifarCNN(dropout=0.0).cuda().train()
    b.load_state_dict(a.state_dict())
    xs = torch.randn(6, 3, 32, 32, device="cuda").bfloat16()
    ys = torch.randint(0, 10, (6,), device="cuda")
    for i in range(6):
        la = a.native_sample_step(xs[i], ys[i], 0.05).clone()
        for p in b.parameters():
            if p.grad is not None:
                p.grad.zero_()
        lb = b.forward_backward(xs[i:i + 1], ys[i:i + 1])
        for p in b.parameters():
            K.sgd_step(p.data.view(-1), p.grad.view(-1), 0.05, 0.0, True)
        assert abs(la.item() - lb.item()) < 1e-5, i
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        assert ((p - q).norm() / q.norm()).item() < 1e-5, n
    # dropout: every replay draws a new mask (the seed counter lives on the device and the graph bumps it)
    c = CifarCNN(dropout=0.5).cuda().train()
    l = [c.native_sample_step(xs[0], ys[0], 0.0).item() for _ in range(4)]
    assert len(set(l)) > 1
    # StochasticGradient picks the graph path up by itself and the training error falls
    torch.manual_seed(1)
    m = CifarCNN(dropout=0.0).cuda().train()
    t = torch.randint(0, 10, (64,), device="cuda")
    pat = torch.randn(10, 3, 32, 32, device="cuda")
    x = (0.3 * torch.randn(64, 3, 32, 32, device="cuda") + pat[t]).bfloat16()
    tr = nn.StochasticGradient(m, nn.CrossEntropyCriterion())
    tr.learningRate, tr.maxIteration, tr.verbose, tr.shuffleIndices = 0.02, 4, False, False
    errs = []
    tr.hookIteration = lambda s, it, e: errs.append(e)
    tr.train((x, t))
    assert "_sample_graph" in m.__dict__
    assert errs[-1] < 0.85 * errs[0], errs
