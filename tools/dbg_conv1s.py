"""Diagnose / time csrc/conv1_sliding.cu on one GPU: per-tap one-hot checks (which K chunk of the sliding operand is
wrong, if any), full random check, CUDA-event timing at B = 4096 against the patch-row kernels."""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tadnn.ops import cnn as K  # noqa: E402


def ref(x, w, b):
    r = lambda t: t.bfloat16().float()
    y = F.conv2d(r(x), r(w).view(64, 3, 5, 5), r(b)).bfloat16().float()
    return F.relu(F.max_pool2d(y, 3, 3)).permute(0, 2, 3, 1).contiguous()


def timeit(fn, n=20):
    flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    torch.manual_seed(0)
    dev = "cuda"
    B = 3
    x = (torch.rand(B, 3, 32, 32, device=dev) + 0.5).bfloat16()
    xq = K.nhwc4_pack(x)
    bad = []
    for kh in range(5):
        for kw in range(5):
            for c in range(3):
                w = torch.zeros(64, 75, device=dev)
                w[:, (c * 5 + kh) * 5 + kw] = 1.0
                w[1::2] *= 0.5
                b = torch.zeros(64, device=dev)
                p1, _ = K.conv1s_pool_forward(xq, K.conv1s_weight_pack(w, b))
                e = (p1.float() - ref(x.float(), w, b)).abs().max().item()
                if e > 1e-3:
                    bad.append((kh, kw, c, round(e, 4)))
    print("one-hot taps wrong:", bad if bad else "none")
    w = torch.zeros(64, 75, device=dev)
    b = torch.linspace(0.1, 1.0, 64, device=dev)
    p1, _ = K.conv1s_pool_forward(xq, K.conv1s_weight_pack(w, b))
    print("bias-only max err:", (p1.float() - ref(x.float(), w, b)).abs().max().item())
    for Bn in (1, 2, 149, 1000):
        x = torch.randn(Bn, 3, 32, 32, device=dev).bfloat16()
        w = torch.randn(64, 75, device=dev) * 0.2
        b = torch.randn(64, device=dev) * 0.1
        p1, idx1 = K.conv1s_pool_forward(K.nhwc4_pack(x), K.conv1s_weight_pack(w, b))
        r = ref(x.float(), w, b)
        err = (p1.float() - r).abs()
        print(f"B={Bn}: max err {err.max().item():.4f} frac>1e-6 {(err > 1e-6).float().mean().item():.5f} "
              f"by pooled row {[round(v, 3) for v in err.amax(dim=(0, 2, 3)).tolist()]}")
    if hasattr(K, "conv1s_pool_backward"):
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
        from test_gpu_conv1_sliding import _ref_backward
        for Bn, rows in ((1, 9), (3, 7), (200, 7)):
            x = torch.randn(Bn, 3, 32, 32, device=dev).bfloat16()
            w = torch.randn(64, 75, device=dev) * 0.2
            b = torch.randn(64, device=dev) * 0.1
            xq = K.nhwc4_pack(x)
            p1, idx1 = K.conv1s_pool_forward(xq, K.conv1s_weight_pack(w, b))
            dp1 = torch.randn(Bn, 9, 9, 64, device=dev).bfloat16()
            gw, gb = torch.zeros(64, 75, device=dev), torch.zeros(64, device=dev)
            K.conv1s_pool_backward(xq, dp1, idx1, gw, gb, rows=rows)
            gw_ref, gb_ref = _ref_backward(x, dp1, idx1, rows)
            e = (gw - gw_ref).view(64, 3, 5, 5)
            print(f"bwd B={Bn} rows={rows}: rel gw {((gw - gw_ref).norm() / gw_ref.norm()).item():.5f} rel gb {((gb - gb_ref).norm() / gb_ref.norm()).item():.5f}"
                  f" | err by kh {[round(v, 3) for v in e.abs().amax(dim=(0, 1, 3)).tolist()]} by kw {[round(v, 3) for v in e.abs().amax(dim=(0, 1, 2)).tolist()]}"
                  f" by c {[round(v, 3) for v in e.abs().amax(dim=(0, 2, 3)).tolist()]} ref max {gw_ref.abs().max().item():.2f}")
    Bn = 4096
    x = torch.randn(Bn, 3, 32, 32, device=dev).bfloat16()
    xu = torch.randint(0, 256, (Bn, 3, 32, 32), device=dev, dtype=torch.uint8)
    w = torch.randn(64, 75, device=dev) * 0.2
    b = torch.randn(64, device=dev) * 0.1
    xq = K.nhwc4_pack(x)
    wx = K.conv1s_weight_pack(w, b)
    wp = K.conv1_weight_pack(w)
    print("conv1s fwd  us (median, min):", timeit(lambda: K.conv1s_pool_forward(xq, wx)))
    print("conv1  fwd  us (median, min):", timeit(lambda: K.conv1_pool_forward(x, wp, b)))
    print("nhwc4_pack u8 us:", timeit(lambda: K.nhwc4_pack(xu, [4 / 255] * 3, [-2.0] * 3)))
    print("u8_to_bf16    us:", timeit(lambda: K.u8_to_bf16(xu, 1024, 3, [4 / 255] * 3, [-2.0] * 3)))
    if hasattr(K, "conv1s_pool_backward"):
        p1, idx1 = K.conv1s_pool_forward(xq, wx)
        dp1 = torch.zeros(Bn, 9, 9, 64, device=dev, dtype=torch.bfloat16)
        dp1[:, :7, :7] = torch.randn(Bn, 7, 7, 64, device=dev).bfloat16()
        gw, gb = torch.zeros(64, 75, device=dev), torch.zeros(64, device=dev)
        print("conv1s bwd  us (median, min):", timeit(lambda: K.conv1s_pool_backward(xq, dp1, idx1, gw, gb, rows=7)))
        p_old, i_old = K.conv1_pool_forward(x, wp, b)
        print("conv1  bwd  us (median, min):", timeit(lambda: K.conv1_pool_backward_bias(x, dp1, p_old, i_old, gw, gb)))


if __name__ == "__main__":
    main()
