"""The fused second stage (csrc/conv2_fused.cu), the re-tiled classifier head and the bias-gradient
side output of conv1's backward, each against a plain PyTorch fp32 statement of the same op."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _pack(w1, w2):
    from tadnn.ops import cnn as K
    dev = w2.device
    w1p = torch.empty((64, 80), device=dev, dtype=torch.bfloat16)
    w2p = torch.empty((64, 1600), device=dev, dtype=torch.bfloat16)
    Tt = torch.empty((576, 4032), device=dev, dtype=torch.bfloat16)
    K.cnn_pack_shadows(w1, w2, w1p, w2p, Tt)
    return w1p, w2p, Tt


def test_pack_shadows_matches_the_separate_pack_kernels():
    from tadnn.ops import cnn as K
    torch.manual_seed(0)
    w1, w2 = torch.randn(64, 75, device="cuda"), torch.randn(64, 1600, device="cuda")
    w1p, w2p, Tt = _pack(w1, w2)
    assert torch.equal(w1p, K.conv1_weight_pack(w1))
    assert torch.equal(w2p, K.conv_weight_pack(w2, 64, 5, 5, 1600))
    # Tt[(pos, o), (fy, fx, c)] = W[o, c, fy - py, fx - px]
    W = w2.view(64, 64, 5, 5).bfloat16()
    ref = torch.zeros(3, 3, 64, 7, 9, 64, device="cuda", dtype=torch.bfloat16)
    for py in range(3):
        for px in range(3):
            ref[py, px, :, py:py + 5, px:px + 5, :] = W.permute(0, 2, 3, 1)
    assert torch.equal(Tt, ref.reshape(576, 4032))


@pytest.mark.parametrize("B", [1, 5, 7, 133])
def test_conv2_pool_forward_matches_pytorch(B):
    from tadnn.ops import cnn as K
    torch.manual_seed(B)
    p1 = torch.relu(torch.randn(B, 9, 9, 64, device="cuda")).bfloat16()
    w2 = torch.randn(64, 1600, device="cuda") * 0.05
    b2 = torch.randn(64, device="cuda") * 0.1
    _, w2p, _ = _pack(torch.zeros(64, 75, device="cuda"), w2)
    p2, idx2, a2 = K.conv2_pool_forward(p1, w2p, b2, store_a2=True)
    ref = F.relu(F.conv2d(p1.float().permute(0, 3, 1, 2), w2.view(64, 64, 5, 5).bfloat16().float(), b2))  # [B,64,5,5]
    ref_a2 = ref.permute(0, 2, 3, 1)
    scale = max(1.0, ref.abs().max().item())
    assert (a2.float() - ref_a2).abs().max().item() < 0.02 * scale
    # pooling is checked against the kernel's own (bf16) un-pooled output: exact values, first arg-max
    win = a2[:, :3, :3, :].float().reshape(B, 9, 64)
    mx, am = win.max(dim=1)
    assert torch.equal(p2.float(), mx)
    first = (win == mx.unsqueeze(1)).float().argmax(dim=1)  # argmax of a 0/1 mask = first maximal position
    assert torch.equal(idx2.long(), first)
    p2b, idx2b = K.conv2_pool_forward(p1, w2p, b2)  # without the optional store
    assert torch.equal(p2b, p2) and torch.equal(idx2b, idx2)


@pytest.mark.parametrize("B", [1, 7, 8, 300])
def test_conv2_wgrad_matches_pytorch(B):
    from tadnn.ops import cnn as K
    torch.manual_seed(B)
    p1 = torch.relu(torch.randn(B, 9, 9, 64, device="cuda")).bfloat16()
    da2w = (torch.randn(B * 9, 64, device="cuda") * (torch.rand(B * 9, 64, device="cuda") < 0.2)).bfloat16()
    dwp = torch.zeros(64, 1600, device="cuda")
    K.conv2_wgrad(p1, da2w, dwp)
    K.conv2_wgrad(p1, da2w, dwp)  # accumulates
    g = da2w.float().view(B, 3, 3, 64)
    x = p1.float()
    ref = torch.zeros(64, 25, 64, device="cuda")
    for kh in range(5):
        for kw in range(5):
            ref[:, kh * 5 + kw, :] = torch.einsum("nyxo,nyxc->oc", g, x[:, kh:kh + 3, kw:kw + 3, :])
    ref = 2 * ref.reshape(64, 1600)
    assert ((dwp - ref).norm() / ref.norm()).item() < 2e-3
    gw = torch.ones(64, 1600, device="cuda")
    K.conv2_wgrad_unpack(dwp, gw)
    assert float(dwp.abs().sum()) == 0.0
    ref_t7 = ref.view(64, 25, 64).permute(0, 2, 1).reshape(64, 1600)
    assert ((gw - 1 - ref_t7).norm() / ref_t7.norm()).item() < 2e-3


def test_toeplitz_dgrad_gemm_is_the_transposed_convolution():
    from tadnn.ops import cnn as K
    torch.manual_seed(0)
    B = 70
    w2 = torch.randn(64, 1600, device="cuda") * 0.05
    _, _, Tt = _pack(torch.zeros(64, 75, device="cuda"), w2)
    da2w = torch.randn(B * 9, 64, device="cuda").bfloat16()
    dp1 = torch.zeros(B, 9, 9, 64, device="cuda", dtype=torch.bfloat16)
    K.gemm_bf16(da2w.view(B, 576), Tt, b_mn_major=True, out=dp1.view(B, 5184)[:, :4032])
    g = torch.zeros(B, 64, 5, 5, device="cuda")
    g[:, :, :3, :3] = da2w.float().view(B, 3, 3, 64).permute(0, 3, 1, 2)
    ref = F.conv_transpose2d(g, w2.view(64, 64, 5, 5).bfloat16().float()).permute(0, 2, 3, 1)  # [B,9,9,64]
    assert (dp1.float() - ref).abs().max().item() < 0.03 * ref.abs().max().item()
    assert float(dp1[:, 7:].abs().sum()) == 0.0 and float(dp1[:, :, 7:].abs().sum()) == 0.0


def _head(B, D, H, NC, p2, w1, b1, w2, b2, tgt, drop=0.0, seed=1):
    from tadnn.ops import cnn as K
    h = K.HeadArgs()
    dev = p2.device
    hidden = torch.empty(B, H, device=dev)
    logp = torch.empty(B, NC, device=dev)
    loss = torch.empty((), device=dev)
    h.p2, h.w1, h.b1, h.w2, h.b2, h.target = p2.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), tgt.data_ptr()
    h.B, h.D, h.H, h.NC, h.dropout_p, h.seed = B, D, H, NC, drop, seed
    h.hidden, h.logp, h.loss = hidden.data_ptr(), logp.data_ptr(), loss.data_ptr()
    return h, hidden, logp, loss


@pytest.mark.parametrize("B,D,H,NC", [(1, 64, 100, 10), (77, 64, 100, 10), (4096, 64, 100, 10), (50, 48, 37, 5), (33, 64, 128, 32)])
def test_head_forward_backward_match_pytorch(B, D, H, NC):
    from tadnn import _native
    from tadnn.ops import cnn as K
    torch.manual_seed(B + D)
    p2 = torch.relu(torch.randn(B, D, device="cuda")).bfloat16()
    w1 = (torch.randn(H, D, device="cuda") * 0.2).requires_grad_(True)
    b1 = (torch.randn(H, device="cuda") * 0.1).requires_grad_(True)
    w2 = (torch.randn(NC, H, device="cuda") * 0.2).requires_grad_(True)
    b2 = (torch.randn(NC, device="cuda") * 0.1).requires_grad_(True)
    tgt = torch.randint(0, NC, (B,), device="cuda")
    h, hidden, logp, loss = _head(B, D, H, NC, p2, w1, b1, w2, b2, tgt)
    lib = _native.lib()
    ws = K.head_workspace(p2.device)
    s = C.c_void_p(_native.stream_ptr())
    _native.check(lib.tadnn_head_forward(C.byref(h), C.c_void_p(ws.data_ptr()), s), "fwd")
    x = p2.float().requires_grad_(True)
    ref_logp = F.log_softmax(F.linear(F.relu(F.linear(x, w1, b1)), w2, b2), dim=-1)
    ref_loss = F.nll_loss(ref_logp, tgt)
    assert (logp - ref_logp.detach()).abs().max().item() < 1e-4
    assert abs(loss.item() - ref_loss.item()) < 1e-5 * max(1.0, abs(ref_loss.item())) + 1e-6
    ref_loss.backward()
    g = [torch.ones(H, D, device="cuda"), torch.ones(H, device="cuda"), torch.ones(NC, H, device="cuda"), torch.ones(NC, device="cuda")]
    dp2 = torch.empty(B, D, device="cuda", dtype=torch.bfloat16)
    h.gw1, h.gb1, h.gw2, h.gb2, h.dp2, h.grad_scale = g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), g[3].data_ptr(), dp2.data_ptr(), 1.0
    if D % 2 == 0:
        idx = torch.randint(0, 9, (B, D), device="cuda", dtype=torch.uint8)
        dpre = torch.full((B * 9, D), 7.0, device="cuda", dtype=torch.bfloat16)
        gpre = torch.ones(D, device="cuda")
        h.pool_idx, h.dpre, h.gbias_pre = idx.data_ptr(), dpre.data_ptr(), gpre.data_ptr()
    _native.check(lib.tadnn_head_backward(C.byref(h), s), "bwd")
    for got, ref in zip(g, (w1.grad, b1.grad, w2.grad, b2.grad)):
        assert ((got - 1 - ref).norm() / (ref.norm() + 1e-12)).item() < 1e-4
    assert (dp2.float() - x.grad).abs().max().item() < 1e-2 * x.grad.abs().max().item() + 1e-8
    if D % 2 == 0:
        live = dp2.float() * (p2.float() > 0)
        want = torch.zeros(B, 9, D, device="cuda")
        want.scatter_(1, idx.long().unsqueeze(1), live.unsqueeze(1))
        assert torch.equal(dpre.float().view(B, 9, D), want)
        assert torch.allclose(gpre - 1, live.sum(0), rtol=1e-4, atol=1e-5)


def test_head_dropout_mask_is_the_same_in_both_directions():
    from tadnn import _native
    from tadnn.ops import cnn as K
    torch.manual_seed(3)
    B, D, H, NC = 64, 64, 100, 10
    p2 = torch.relu(torch.randn(B, D, device="cuda")).bfloat16() + 1
    w1, b1 = torch.randn(H, D, device="cuda") * 0.2, torch.zeros(H, device="cuda")
    w2, b2 = torch.randn(NC, H, device="cuda") * 0.2, torch.zeros(NC, device="cuda")
    tgt = torch.randint(0, NC, (B,), device="cuda")
    h, hidden, logp, loss = _head(B, D, H, NC, p2, w1, b1, w2, b2, tgt, drop=0.5, seed=99)
    lib, s = _native.lib(), C.c_void_p(_native.stream_ptr())
    _native.check(lib.tadnn_head_forward(C.byref(h), C.c_void_p(K.head_workspace(p2.device).data_ptr()), s), "fwd")
    g = [torch.zeros(H, D, device="cuda"), torch.zeros(H, device="cuda"), torch.zeros(NC, H, device="cuda"), torch.zeros(NC, device="cuda")]
    dp2 = torch.empty(B, D, device="cuda", dtype=torch.bfloat16)
    h.gw1, h.gb1, h.gw2, h.gb2, h.dp2, h.grad_scale = g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), g[3].data_ptr(), dp2.data_ptr(), 1.0
    _native.check(lib.tadnn_head_backward(C.byref(h), s), "bwd")
    dropped = dp2.float() == 0  # a dropped input receives no gradient; about half of them
    assert 0.35 < dropped.float().mean().item() < 0.65
    # fc1's weight gradient has exactly zero columns contribution from dropped inputs: check one sample's pattern
    x = (p2.float() * (~dropped) * 2.0).requires_grad_(True)  # the same mask, applied by hand
    ref = F.nll_loss(F.log_softmax(F.linear(F.relu(F.linear(x, w1, b1)), w2, b2), -1), tgt)
    assert abs(ref.item() - loss.item()) < 1e-4


@pytest.mark.parametrize("B", [3, 130])
def test_conv1_backward_bias_side_output(B):
    from tadnn.ops import cnn as K
    torch.manual_seed(B)
    x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
    w = torch.randn(64, 75, device="cuda") * 0.2
    b = torch.randn(64, device="cuda") * 0.1
    p1, idx1 = K.conv1_pool_forward(x, K.conv1_weight_pack(w), b)
    dp1 = torch.randn(B, 9, 9, 64, device="cuda").bfloat16()
    gw, gw_ref = torch.zeros(64, 75, device="cuda"), torch.zeros(64, 75, device="cuda")
    gb, gb_ref = torch.ones(64, device="cuda"), torch.ones(64, device="cuda")
    K.conv1_pool_backward_bias(x, dp1, p1, idx1, gw, gb)
    K.conv1_pool_backward(x, dp1, p1, idx1, gw_ref)
    K.colsum_relu_mask(dp1, p1, gb_ref)
    assert torch.equal(gw, gw_ref) or ((gw - gw_ref).norm() / gw_ref.norm()).item() < 1e-5
    assert torch.allclose(gb, gb_ref, rtol=1e-4, atol=1e-3)


def test_gemm_kblock_mask_skips_only_structural_zeros():
    """The transposed conv2 as a GEMM against the Toeplitz matrix: with the k-block mask (structural zeros skipped) the result
    is bit-identical to the dense product."""
    from tadnn.models.cifar_cnn import _TT_MASK
    from tadnn.ops import cnn as K
    torch.manual_seed(3)
    w1 = torch.randn(64, 75, device="cuda")
    w2 = torch.randn(64, 1600, device="cuda") * 0.05
    w1p = torch.empty(64, 80, device="cuda", dtype=torch.bfloat16)
    w2p = torch.empty(64, 1600, device="cuda", dtype=torch.bfloat16)
    Tt = torch.empty(576, 4032, device="cuda", dtype=torch.bfloat16)
    K.cnn_pack_shadows(w1, w2, w1p, w2p, Tt)
    assert sum(bin(m).count("1") for m in _TT_MASK) < 0.7 * 9 * len(_TT_MASK)
    for B in (5, 300, 1000):
        g = torch.randn(B, 576, device="cuda").bfloat16()
        dense = K.gemm_bf16(g, Tt, b_mn_major=True, tile_n=256)
        sparse = K.gemm_bf16(g, Tt, b_mn_major=True, tile_n=256, kblock_mask=_TT_MASK)
        assert torch.equal(dense, sparse)
        ref = g.float() @ Tt.float()
        assert ((sparse.float() - ref).norm() / ref.norm()).item() < 1e-2
