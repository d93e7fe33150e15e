"""csrc/conv1_sliding.cu: the first conv + ReLU + max-pool stage with the staged NHWC4 image as the tcgen05 operand
(no im2col rows), against PyTorch fp32 on the same bf16-rounded operands and against the patch-row kernels of
csrc/conv1_fused.cu."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref_forward(x, w, b):
    """fp32 reference on bf16-rounded operands; conv output rounded to bf16 before pooling, like the kernels."""
    r = lambda t: t.bfloat16().float()
    y = F.conv2d(r(x), r(w).view(64, 3, 5, 5), r(b)).bfloat16().float()
    p, idx = F.max_pool2d(y, 3, 3, return_indices=True)     # [B,64,9,9]; idx = flat index in the 28x28 plane
    oy, ox = idx // 28, idx % 28
    py = torch.arange(9, device=x.device).view(1, 1, 9, 1)
    px = torch.arange(9, device=x.device).view(1, 1, 1, 9)
    pos = (oy - 3 * py) * 3 + (ox - 3 * px)
    return F.relu(p).permute(0, 2, 3, 1).contiguous(), pos.permute(0, 2, 3, 1).contiguous(), y


def test_nhwc4_pack():
    from tadnn.ops import cnn as K
    torch.manual_seed(0)
    xu = torch.randint(0, 256, (5, 3, 32, 32), device="cuda", dtype=torch.uint8)
    sc, sh = [4 / 255, 2 / 255, 1 / 255], [-2.0, -1.0, 0.5]
    q = K.nhwc4_pack(xu, sc, sh)
    ref = xu.float() * torch.tensor(sc, device="cuda").view(1, 3, 1, 1) + torch.tensor(sh, device="cuda").view(1, 3, 1, 1)
    assert (q[..., :3].float() - ref.permute(0, 2, 3, 1)).abs().max().item() <= 2 ** -7
    assert torch.all(q[..., 3] == 1)
    for dt in (torch.bfloat16, torch.float32):
        x = torch.randn(3, 3, 32, 32, device="cuda").to(dt)
        q = K.nhwc4_pack(x)
        assert torch.equal(q[..., :3], x.bfloat16().permute(0, 2, 3, 1)) and torch.all(q[..., 3] == 1)


@pytest.mark.parametrize("B", [1, 3, 150, 700])
def test_conv1s_forward_matches_pytorch(B):
    from tadnn.ops import cnn as K
    torch.manual_seed(B)
    x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
    w = torch.randn(64, 75, device="cuda") * 0.2
    b = torch.randn(64, device="cuda") * 0.1
    p1, idx1 = K.conv1s_pool_forward(K.nhwc4_pack(x), K.conv1s_weight_pack(w, b))
    p_ref, pos_ref, _ = _ref_forward(x.float(), w, b)
    err = (p1.float() - p_ref).abs()
    assert err.max().item() <= 0.02 * max(1.0, p_ref.abs().max().item()), err.max().item()
    assert (err > 1e-6).float().mean().item() < 0.02     # fp32 summation order flips a bf16 rounding now and then
    alive = p_ref > 0
    # idx1 = 8 - window position, 15 = dead window (pooled ReLU output 0)
    assert torch.all(idx1[p1 > 0] <= 8) and (idx1[p1 == 0] != 15).float().mean().item() < 1e-4
    pos = 8 - idx1.long()
    same = (pos == pos_ref) | ~alive
    assert same.float().mean().item() > 0.995
    # and against the patch-row kernel (same rounding points)
    p_old, idx_old = K.conv1_pool_forward(x, K.conv1_weight_pack(w), b)
    assert (p1.float() - p_old.float()).abs().max().item() <= 0.02 * max(1.0, p_ref.abs().max().item())
    assert ((pos == idx_old.long()) | (p_old <= 0)).float().mean().item() > 0.995


def _ref_backward(x, dp1, idx1, rows):
    """dW, db of conv1 for d(conv out) rebuilt from (dp1, idx1) in plain PyTorch fp32."""
    B = x.shape[0]
    dy = torch.zeros(B, 64, 28, 28, device=x.device)
    g = dp1.float().clone()
    g[:, rows:] = 0
    pos = 8 - idx1.long()
    alive = idx1 != 15
    py = torch.arange(9, device=x.device).view(1, 9, 1, 1).expand(B, 9, 9, 64)
    px = torch.arange(9, device=x.device).view(1, 1, 9, 1).expand(B, 9, 9, 64)
    oy, ox = 3 * py + pos // 3, 3 * px + pos % 3
    n = torch.arange(B, device=x.device).view(B, 1, 1, 1).expand(B, 9, 9, 64)
    c = torch.arange(64, device=x.device).view(1, 1, 1, 64).expand(B, 9, 9, 64)
    dy[n[alive], c[alive], oy[alive], ox[alive]] = g[alive]
    gw = torch.nn.grad.conv2d_weight(x.float(), (64, 3, 5, 5), dy)
    return gw.reshape(64, 75), dy.sum(dim=(0, 2, 3))


@pytest.mark.parametrize("B,rows", [(1, 9), (2, 7), (150, 7), (333, 9)])
def test_conv1s_backward_matches_pytorch(B, rows):
    from tadnn.ops import cnn as K
    torch.manual_seed(B)
    x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
    w = torch.randn(64, 75, device="cuda") * 0.2
    b = torch.randn(64, device="cuda") * 0.1
    xq = K.nhwc4_pack(x)
    p1, idx1 = K.conv1s_pool_forward(xq, K.conv1s_weight_pack(w, b))
    dp1 = torch.randn(B, 9, 9, 64, device="cuda").bfloat16()
    gw = torch.full((64, 75), 0.5, device="cuda")
    gb = torch.full((64,), -0.25, device="cuda")
    K.conv1s_pool_backward(xq, dp1, idx1, gw, gb, rows=rows)
    gw_ref, gb_ref = _ref_backward(x, dp1, idx1, rows)
    assert ((gw - 0.5 - gw_ref).norm() / gw_ref.norm()).item() < 2e-3
    assert ((gb + 0.25 - gb_ref).norm() / gb_ref.norm()).item() < 2e-3
    # accumulates: a second call doubles the increment
    K.conv1s_pool_backward(xq, dp1, idx1, gw, gb, rows=rows)
    assert ((gw - 0.5 - 2 * gw_ref).norm() / gw_ref.norm()).item() < 4e-3
