"""GEMV-style classifier-layer kernels (csrc/smalln.cu) and the GPU dataset preprocessing
(csrc/preprocess.cu) against plain PyTorch fp32 references of the same ops."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(200, 10, 256), (4096, 10, 256), (37, 16, 64), (513, 3, 2048), (8, 1, 8)])
def test_smalln_fwd_dgrad_wgrad(M, N, K):
    from tadnn.ops import smalln
    torch.manual_seed(M + N + K)
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    b = torch.randn(N, device="cuda")
    pitch = (N + 7) // 8 * 8
    g_store = torch.randn(M, pitch, device="cuda").bfloat16()
    g = g_store[:, :N]  # padded row pitch, as FusedMPMLP's replicated gradient buffer
    aux = torch.tanh(torch.randn(M, K, device="cuda")).bfloat16()

    y = smalln.linear_smalln_fwd(x, w, bias=b)
    ref = x.float() @ w.float().t() + b
    assert (y - ref).abs().max().item() < 1e-3 * max(1.0, ref.abs().max().item())
    y2 = smalln.linear_smalln_fwd(x, w)  # no bias, second call in a row
    assert (y2 - (ref - b)).abs().max().item() < 1e-3 * max(1.0, ref.abs().max().item())

    dx = smalln.linear_smalln_dgrad(g, w)
    dref = g.float() @ w.float()
    assert (dx.float() - dref).abs().max().item() < 1e-2 * max(1.0, dref.abs().max().item())
    dxa = smalln.linear_smalln_dgrad(g, w, aux_tanh_out=aux)
    drefa = dref * (1 - aux.float() ** 2)
    assert (dxa.float() - drefa).abs().max().item() < 1e-2 * max(1.0, drefa.abs().max().item())

    dw, db = smalln.linear_smalln_wgrad(g, x)
    wref = g.float().t() @ x.float()
    bref = g.float().sum(0)
    assert ((dw - wref).norm() / wref.norm()).item() < 1e-4
    assert (db - bref).abs().max().item() < 1e-3 * max(1.0, bref.abs().max().item())
    smalln.linear_smalln_wgrad(g, x, dw=dw, db=db, accumulate=True)
    assert ((dw - 2 * wref).norm() / wref.norm()).item() < 2e-4
    assert (db - 2 * bref).abs().max().item() < 2e-3 * max(1.0, bref.abs().max().item())


def test_smalln_strided_input_view():
    """The first layer of FusedMPMLP reads a column slice of the input in place (narrowInput)."""
    from tadnn.ops import smalln
    xf = torch.randn(300, 512, device="cuda").bfloat16()
    x = xf[:, 128:256]
    w = torch.randn(10, 128, device="cuda").bfloat16()
    y = smalln.linear_smalln_fwd(x, w)
    ref = x.float() @ w.float().t()
    assert (y - ref).abs().max().item() < 1e-3 * ref.abs().max().item()


def test_smalln_rejects_wide_or_unaligned():
    from tadnn import _native
    from tadnn.ops import smalln
    assert not smalln.supported(17, 64) and not smalln.supported(10, 60)
    x = torch.zeros(4, 64, device="cuda", dtype=torch.bfloat16)
    w = torch.zeros(17, 64, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(4, 17, device="cuda")
    import ctypes as C
    rc = _native.lib().tadnn_linear_smalln_fwd(C.c_void_p(x.data_ptr()), 64, C.c_void_p(w.data_ptr()), 64, None,
                                               C.c_void_p(out.data_ptr()), 17, 4, 17, 64, None)
    assert rc != 0 and b"1..16" in _native.lib().tadnn_last_error()


@pytest.mark.parametrize("shape", [(64, 32, 32), (5, 24, 40)])
def test_preprocess_cuda_matches_torch(shape):
    from tadnn.utils import preprocess as P
    n, H, W = shape
    torch.manual_seed(1)
    train = torch.rand(n, 3, H, W, device="cuda")
    test = torch.rand(max(n // 4, 1), 3, H, W, device="cuda")
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False  # the reference below must be a true fp32 convolution
    try:
        tr_ref, te_ref = P.preprocess(train, test)
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
    tr, te = P.preprocess_cuda(train, test)
    for got, ref in ((tr, tr_ref), (te, te_ref)):
        assert got.shape == ref.shape
        assert (got - ref).abs().max().item() < 2e-3 * max(1.0, ref.abs().max().item())
    trn, _ = P.preprocess_cuda(train, None, reference_sign=True)  # Q10: the reference divides by -std
    assert torch.allclose(trn[:, 1:], -tr[:, 1:], atol=1e-5) and torch.allclose(trn[:, 0], tr[:, 0])
    trb, _ = P.preprocess_cuda(train, None, out_dtype=torch.bfloat16)
    assert trb.dtype == torch.bfloat16 and (trb.float() - tr).abs().max().item() < 2e-2 * max(1.0, tr.abs().max().item())


def test_preprocess_cuda_flat_image_is_zero():
    from tadnn.utils import preprocess as P
    flat = torch.full((2, 3, 32, 32), 0.5, device="cuda")
    out = P.yuv_lcn_cuda(flat, (0.0, 1.0, 0.0, 1.0))
    assert out[:, 0].abs().max().item() < 1e-2


@pytest.mark.parametrize("shape", [(64, 3, 32, 32), (5, 3, 7, 9), (33, 1, 16, 16)])
def test_u8_to_bf16_matches_torch(shape):
    from tadnn.ops.cnn import u8_to_bf16
    torch.manual_seed(0)
    x = torch.randint(0, 256, shape, dtype=torch.uint8, device="cuda")
    C_ = shape[1]
    scale = [0.01 * (c + 1) for c in range(C_)]
    shift = [-1.0 + 0.25 * c for c in range(C_)]
    y = u8_to_bf16(x, shape[2] * shape[3], C_, scale, shift)
    ref = x.float() * torch.tensor(scale, device="cuda").view(1, C_, 1, 1) + torch.tensor(shift, device="cuda").view(1, C_, 1, 1)
    assert y.dtype == torch.bfloat16 and y.shape == x.shape
    assert (y.float() - ref).abs().max().item() <= 2 ** -7 * max(1.0, ref.abs().max().item())  # one bf16 ulp (fma vs mul+add)
