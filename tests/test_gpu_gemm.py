"""tcgen05 GEMM numerics vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn, bias, act, alpha):
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    y = alpha * (A @ B.t())
    if bias is not None:
        y = y + bias
    if act == "tanh":
        y = torch.tanh(y)
    elif act == "relu":
        y = torch.relu(y)
    return y


def _check(M, N, K, a_mn, b_mn, out_dtype=torch.bfloat16, bias=False, act=None, alpha=1.0,
           split_k=0, accumulate=False, tile_n=0):
    from tadnn.ops.gemm import gemm_bf16
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda", generator=g).bfloat16()
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda", generator=g).bfloat16()
    bv = torch.randn(N, device="cuda", generator=g) if bias else None
    scale = alpha / (K ** 0.5) if act is None else alpha * 0.2 / (K ** 0.5)
    ref = _ref(a, b, a_mn, b_mn, bv, act, scale)
    out = None
    if accumulate:
        out = torch.ones((M, N), device="cuda", dtype=torch.float32)
        ref = ref + 1.0
    y = gemm_bf16(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out=out, out_dtype=out_dtype, bias=bv,
                  act=act, alpha=scale, split_k=split_k, accumulate=accumulate, tile_n=tile_n)
    torch.cuda.synchronize()
    tol = 2e-2 if out_dtype == torch.bfloat16 else 2e-3
    err = (y.float() - ref).abs().max().item()
    assert err < tol * max(1.0, ref.abs().max().item()), f"max err {err} (M={M},N={N},K={K},a_mn={a_mn},b_mn={b_mn})"


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 256), (1000, 2048, 1024), (4096, 4096, 4096),
                                   (300, 136, 72), (64, 64, 1600)])
def test_gemm_tn(M, N, K):
    _check(M, N, K, False, False)


@pytest.mark.parametrize("M,N,K", [(256, 512, 256), (1000, 1024, 2048), (200, 136, 72)])
def test_gemm_dgrad_layout(M, N, K):
    _check(M, N, K, False, True)


@pytest.mark.parametrize("M,N,K", [(256, 512, 256), (2048, 1024, 1000), (64, 1600, 4000), (136, 200, 520)])
def test_gemm_wgrad_layout(M, N, K):
    _check(M, N, K, True, True, out_dtype=torch.float32)


def test_gemm_a_mn_b_k():
    _check(256, 256, 512, True, False)


@pytest.mark.parametrize("tile_n", [64, 128, 256])
def test_gemm_tiles(tile_n):
    _check(512, 512, 512, False, False, tile_n=tile_n)


def test_gemm_bias_tanh():
    _check(512, 2048, 1024, False, False, bias=True, act="tanh")


def test_gemm_bias_relu_fp32():
    _check(384, 64, 80, False, False, out_dtype=torch.float32, bias=True, act="relu")


def test_gemm_splitk_and_accumulate():
    _check(64, 1600, 8192, True, True, out_dtype=torch.float32, split_k=16)
    _check(128, 256, 1024, False, False, out_dtype=torch.float32, accumulate=True)
    _check(128, 256, 4096, False, False, out_dtype=torch.float32, accumulate=True, split_k=4)


def test_gemm_perf_smoke():
    """Not a benchmark: prints achieved TFLOP/s at 8192^3 for the run log."""
    from tadnn.ops.gemm import gemm_bf16
    M = N = K = 8192
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        gemm_bf16(a, b, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        gemm_bf16(a, b, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"\n[gemm 8192^3] {ms:.3f} ms  {2 * M * N * K / ms / 1e9:.1f} TFLOP/s")
    ref = a[:256].float() @ b.float().t()
    assert (out[:256].float() - ref).abs().max().item() < 2.0


def test_gemm_allgather_push_epilogue():
    """n_push: every output tile is also stored to the push targets at a column offset (the fused
    all-gather; on one GPU the "peers" are two local buffers)."""
    from tadnn.ops.gemm import gemm_bf16
    M, N, K, P, r = 300, 192, 256, 3, 1
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    peers = [torch.zeros(M, P * N, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
    out = gemm_bf16(a, b, bias=bias, act="tanh", alpha=0.05, push=[p.data_ptr() for p in peers], push_ldc=P * N,
                    push_col_offset=r * N)
    torch.cuda.synchronize()
    ref = torch.tanh(0.05 * (a.float() @ b.float().t()) + bias)
    assert (out.float() - ref).abs().max().item() < 2e-2
    for p in peers:
        assert torch.equal(p[:, r * N:(r + 1) * N], out)
        assert float(p[:, :r * N].abs().sum()) == 0 and float(p[:, (r + 1) * N:].abs().sum()) == 0


def test_gemm_reduce_scatter_push_epilogue():
    """rs_width: column block q of the partial product goes only to push target q (row-major and
    slab-major staging)."""
    from tadnn.ops.gemm import gemm_bf16
    M, N, K, P = 200, 512, 128, 4
    w = N // P
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    ref = (a.float() @ b.float().t()).bfloat16()
    slots = [torch.zeros(M, w, device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    gemm_bf16(a, b, push=[s.data_ptr() for s in slots], push_ldc=w, rs_width=w, n_rotate=1)
    torch.cuda.synchronize()
    for q, s in enumerate(slots):
        assert (s.float() - ref[:, q * w:(q + 1) * w].float()).abs().max().item() <= 0.01 * ref.float().abs().max().item()
    Mp = (M + 31) // 32 * 32
    slabs = [torch.zeros(Mp * w, device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    gemm_bf16(a, b, push=[s.data_ptr() for s in slabs], push_ldc=w, rs_width=w, rs_slab_major=True)
    torch.cuda.synchronize()
    for q, s in enumerate(slabs):
        t = s.view(Mp // 32, w // 64, 32, 8, 8)            # [rowblk, colslab, row, chunk position, 8 elems]
        rows = torch.arange(32, device="cuda") & 7
        pos = torch.arange(8, device="cuda")
        idx = (pos.view(1, 8) ^ rows.view(32, 1))           # chunk j sits at position j ^ (row & 7)
        un = torch.gather(t, 3, idx.view(1, 1, 32, 8, 1).expand(Mp // 32, w // 64, 32, 8, 8))
        dense = un.permute(0, 2, 1, 3, 4).reshape(Mp, w)[:M]
        assert torch.equal(dense, slots[q])


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 768, 1024), (1000, 520, 200), (300, 136, 72), (4096, 3136, 576)])
def test_gemm_2cta_cluster_kernel(M, N, K):
    """tcgen05.mma.cta_group::2 kernel (clusters of two CTAs) vs fp32 reference, incl. ragged edges."""
    from tadnn.ops.gemm import gemm_bf16_2cta
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = torch.randn(N, K, device="cuda", generator=g).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    s = 1.0 / K ** 0.5
    y = gemm_bf16_2cta(a, b, bias=bias, act="relu", alpha=s)
    torch.cuda.synchronize()
    ref = torch.relu(s * (a.float() @ b.float().t()) + bias)
    assert (y.float() - ref).abs().max().item() < 2e-2 * max(1.0, ref.abs().max().item())
