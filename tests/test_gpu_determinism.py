"""Race smoke test.  compute-sanitizer is closed on the GPU pool this project is developed on
(profiles/sanitizer/compute_sanitizer_closed.txt), so the mbarrier / TMEM / named-barrier pipelines are checked
the way a race shows up in practice: the same launch repeated many times must give the same bits (kernels
without floating-point atomics) or the same values to rounding (kernels that finish with red.global.add)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _pack(w1, w2):
    from tadnn.ops import cnn as K
    w1p = torch.empty((64, 80), device="cuda", dtype=torch.bfloat16)
    w2p = torch.empty((64, 1600), device="cuda", dtype=torch.bfloat16)
    Tt = torch.empty((576, 4032), device="cuda", dtype=torch.bfloat16)
    K.cnn_pack_shadows(w1, w2, w1p, w2p, Tt)
    return w1p, w2p, Tt


def test_forward_pipelines_are_bit_reproducible():
    from tadnn.ops import cnn as K
    from tadnn.ops.gemm import gemm_bf16
    torch.manual_seed(0)
    B = 333
    x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
    w1, b1 = torch.randn(64, 75, device="cuda") * 0.2, torch.randn(64, device="cuda") * 0.1
    w2, b2 = torch.randn(64, 1600, device="cuda") * 0.05, torch.randn(64, device="cuda") * 0.1
    w1p, w2p, _ = _pack(w1, w2)
    a, b = torch.randn(1000, 512, device="cuda").bfloat16(), torch.randn(768, 512, device="cuda").bfloat16()
    ref = None
    for it in range(25):
        p1, idx1 = K.conv1_pool_forward(x, w1p, b1)
        p2, idx2 = K.conv2_pool_forward(p1, w2p, b2)
        y = gemm_bf16(a, b, act="tanh")
        y2 = gemm_bf16(a, b)  # 2-CTA kernel
        cur = (p1, idx1, p2, idx2, y, y2)
        if ref is None:
            ref = [t.clone() for t in cur]
        else:
            for i, (r, c) in enumerate(zip(ref, cur)):
                assert torch.equal(r, c), f"output {i} changed on repetition {it}"


def test_backward_pipelines_are_reproducible_to_rounding():
    from tadnn.ops import cnn as K
    torch.manual_seed(1)
    B = 200
    x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
    w1, b1 = torch.randn(64, 75, device="cuda") * 0.2, torch.randn(64, device="cuda") * 0.1
    w2 = torch.randn(64, 1600, device="cuda") * 0.05
    w1p, w2p, _ = _pack(w1, w2)
    p1, idx1 = K.conv1_pool_forward(x, w1p, b1)
    dp1 = torch.randn(B, 9, 9, 64, device="cuda").bfloat16()
    da2w = torch.randn(B * 9, 64, device="cuda").bfloat16()
    ref = None
    for it in range(15):
        gw, gb = torch.zeros(64, 75, device="cuda"), torch.zeros(64, device="cuda")
        K.conv1_pool_backward_bias(x, dp1, p1, idx1, gw, gb)
        dwp = torch.zeros(64, 1600, device="cuda")
        K.conv2_wgrad(p1, da2w, dwp)
        cur = (gw, gb, dwp)
        if ref is None:
            ref = [t.clone() for t in cur]
        else:
            for i, (r, c) in enumerate(zip(ref, cur)):
                rel = ((r - c).norm() / (r.norm() + 1e-12)).item()
                assert rel < 1e-5, f"output {i} differs by {rel} on repetition {it}"
