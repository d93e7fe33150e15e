"""Host-side pieces of the fused CNN path that need no GPU: the k-block mask of the transposed-conv2 Toeplitz matrix,
the packed-input predicate, and the reference statement of the sliding-operand conv1's pooling order."""
import torch
import torch.nn.functional as F


def _toeplitz(w2: torch.Tensor) -> torch.Tensor:
    """Tt[pos*64 + o, (fy*9 + fx)*64 + c] = w2[o, c, fy - pos // 3, fx - pos % 3] (zero outside the 5x5 filter):
    what csrc/conv2_fused.cu:cnn_pack_shadows_kernel builds, in plain PyTorch."""
    Tt = torch.zeros(9 * 64, 7 * 9 * 64)
    for pos in range(9):
        for fy in range(7):
            for fx in range(9):
                kh, kw = fy - pos // 3, fx - pos % 3
                if 0 <= kh < 5 and 0 <= kw < 5:
                    Tt[pos * 64:(pos + 1) * 64, (fy * 9 + fx) * 64:(fy * 9 + fx + 1) * 64] = w2[:, :, kh, kw]
    return Tt


def test_toeplitz_kblock_mask_marks_exactly_the_nonzero_blocks():
    from tadnn.models.cifar_cnn import _toeplitz_kblock_mask
    torch.manual_seed(0)
    w2 = torch.randn(64, 64, 5, 5) + 3.0          # no accidental zeros
    Tt = _toeplitz(w2)
    for tile_n in (64, 128, 256):
        masks = _toeplitz_kblock_mask(tile_n)
        assert len(masks) == (4032 + tile_n - 1) // tile_n
        for t, m in enumerate(masks):
            cols = Tt[:, t * tile_n:(t + 1) * tile_n]
            for kb in range(9):
                nonzero = bool(cols[kb * 64:(kb + 1) * 64].abs().sum() > 0)
                assert nonzero == bool((m >> kb) & 1), (tile_n, t, kb)
            if tile_n == 256:
                assert m != 0     # tadnn_gemm_bf16 wants at least one k-block per column tile (narrower tiles can be all zero)
    assert sum(bin(m).count("1") for m in _toeplitz_kblock_mask(256)) == 90   # of 144 blocks


def test_packed_input_predicate_and_defaults():
    from tadnn.models.cifar_cnn import CifarCNN, is_packed_input
    assert is_packed_input(torch.zeros(2, 32, 32, 4, dtype=torch.bfloat16))
    assert not is_packed_input(torch.zeros(2, 3, 32, 32, dtype=torch.bfloat16))
    assert not is_packed_input(torch.zeros(2, 32, 32, 4))
    assert CifarCNN().packs_input and not CifarCNN(conv1_kernel="rows").packs_input
    assert not CifarCNN(conv2_mode="pooled").packs_input and not CifarCNN(fused_conv1=False).packs_input


def test_reference_statement_pools_before_rounding_for_the_sliding_conv1():
    """Same pooled VALUES either way (rounding is monotonic); what differs is which of two conv outputs that round to the
    same bf16 receives the gradient: the larger fp32 one (sliding-operand kernels) or the first one (patch-row kernels)."""
    from tadnn.models.cifar_cnn import CifarCNN
    torch.manual_seed(1)
    x = torch.randn(3, 3, 32, 32).bfloat16().float()
    a, b = CifarCNN(dropout=0.0), CifarCNN(dropout=0.0, conv1_kernel="rows")
    b.load_state_dict(a.state_dict())
    la = a.eval().reference_logp(x, emulate_bf16=True)
    lb = b.eval().reference_logp(x, emulate_bf16=True)
    assert torch.allclose(la, lb, atol=1e-6)
    # a constructed tie: two conv outputs 1 ulp(fp32) apart round to the same bf16
    y = torch.zeros(1, 1, 3, 3)
    y[0, 0, 0, 0] = 1.0
    y[0, 0, 2, 2] = 1.0 + 2.0 ** -20
    y.requires_grad_(True)
    F.max_pool2d(y, 3, 3).sum().backward()
    assert y.grad[0, 0, 2, 2] == 1 and y.grad[0, 0, 0, 0] == 0          # pool the fp32 values: the larger one wins
    y2 = y.detach().bfloat16().float().requires_grad_(True)
    F.max_pool2d(y2, 3, 3).sum().backward()
    assert y2.grad[0, 0, 0, 0] == 1                                     # round first: the tie goes to the first position
