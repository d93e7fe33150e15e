"""The stock layers on a GPU run on the tadnn kernels (ops/native_nn.py), not on cuBLAS/cuDNN:
numerics against plain PyTorch fp32 for shapes that are NOT the example's, and a launch-table check
that the native kernels are what ran."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


@pytest.mark.parametrize("B,i,o", [(77, 1000, 333), (256, 256, 10), (5, 24, 7), (512, 2048, 1024), (1, 64, 100)])
def test_linear_native_matches_pytorch(B, i, o):
    from tadnn import _native, nn
    torch.manual_seed(B + i)
    lin = nn.Linear(i, o).cuda()
    x = torch.randn(B, i, device="cuda", requires_grad=True)
    n0 = _native.launch_count()
    y = lin(x)
    g = torch.randn_like(y)
    y.backward(g)
    assert _native.launch_count() > n0, "the native kernels did not run"
    assert y.dtype == torch.float32 and y.shape == (B, o)
    xr = x.detach().clone().requires_grad_(True)
    yr = F.linear(xr, lin.weight.detach(), lin.bias.detach())
    gw, gb = torch.autograd.grad(yr, [xr], g, retain_graph=True)[0], None
    assert _rel(y, yr) < 1e-2
    assert _rel(x.grad, gw) < 1.5e-2
    wr = lin.weight.detach().clone().requires_grad_(True)
    br = lin.bias.detach().clone().requires_grad_(True)
    F.linear(x.detach(), wr, br).backward(g)
    assert _rel(lin.weight.grad, wr.grad) < 1.5e-2
    assert _rel(lin.bias.grad, br.grad) < 1.5e-2


def test_linear_native_accepts_bf16_and_1d_and_3d_inputs():
    from tadnn import nn
    torch.manual_seed(0)
    lin = nn.Linear(64, 32).cuda()
    x = torch.randn(3, 5, 64, device="cuda")
    ref = F.linear(x, lin.weight, lin.bias)
    assert _rel(lin(x), ref) < 1e-2 and lin(x).shape == (3, 5, 32)
    assert _rel(lin(x[0, 0]), ref[0, 0]) < 1e-2 and lin(x[0, 0]).shape == (32,)
    yb = lin(x.bfloat16())
    assert yb.dtype == torch.bfloat16 and _rel(yb, ref) < 2e-2


@pytest.mark.parametrize("B,C,O,k,H", [(4, 16, 32, 3, 48), (3, 3, 64, 5, 32), (2, 64, 64, 5, 9), (5, 8, 16, 1, 7)])
def test_conv_native_matches_pytorch(B, C, O, k, H):
    from tadnn import _native, nn
    torch.manual_seed(C + O)
    conv = nn.SpatialConvolutionMM(C, O, k, k).cuda()
    x = torch.randn(B, C, H, H, device="cuda", requires_grad=True)
    n0 = _native.launch_count()
    y = conv(x)
    assert _native.launch_count() > n0
    assert y.shape == (B, O, H - k + 1, H - k + 1) and y.dtype == torch.bfloat16
    g = torch.randn(y.shape, device="cuda")
    y.backward(g)
    xr = x.detach().clone().requires_grad_(True)
    wr = conv.weight.detach().clone().requires_grad_(True)
    br = conv.bias.detach().clone().requires_grad_(True)
    yr = F.conv2d(xr, wr.view(O, C, k, k), br)
    yr.backward(g)
    assert _rel(y, yr) < 1e-2
    assert _rel(conv.weight.grad, wr.grad) < 2e-2
    assert _rel(conv.bias.grad, br.grad) < 2e-2
    assert _rel(x.grad, xr.grad) < 2e-2


def test_maxpool_native_on_conv_output_and_fallbacks():
    from tadnn import nn
    torch.manual_seed(1)
    conv = nn.SpatialConvolutionMM(8, 16, 3, 3).cuda()
    pool = nn.SpatialMaxPooling(2, 2, 2, 2)
    x = torch.randn(3, 8, 11, 11, device="cuda", requires_grad=True)
    a = conv(x)
    y = pool(a)               # bf16, NHWC memory -> native kernel, no ReLU assumption
    g = torch.randn(y.shape, device="cuda")
    y.backward(g)
    ar = a.detach().float().contiguous().requires_grad_(True)
    yr = F.max_pool2d(ar, 2, 2)
    assert torch.equal(y.float(), yr)
    yr.backward(g.bfloat16().float())
    xr = x.detach().clone().requires_grad_(True)
    F.conv2d(xr, conv.weight.detach().view(16, 8, 3, 3), conv.bias.detach()).backward(ar.grad)
    assert _rel(x.grad, xr.grad) < 2e-2
    # strided / padded pooling and fp32 NCHW inputs take the PyTorch statement
    p2 = nn.SpatialMaxPooling(3, 3, 2, 2)
    z = torch.randn(2, 4, 9, 9, device="cuda")
    assert torch.equal(p2(z), F.max_pool2d(z, 3, 2))


def test_example_cnn_as_a_sequential_of_stock_layers_runs_native():
    """cifar_example/sgd-torchad_nn-cifar.lua:88-118 built from stock layers (what the reference's API is for)."""
    from tadnn import _native
    from tadnn.models.cifar_cnn import CifarCNN
    torch.manual_seed(0)
    m = CifarCNN(dropout=0.0).cuda().eval()
    with torch.no_grad():
        for p in m.parameters():
            p.mul_(2.0)
    seq = m.to_sequential().eval()
    x = torch.randn(64, 3, 32, 32, device="cuda").bfloat16().float()   # bf16-exact inputs: both paths see the same numbers
    t = torch.randint(0, 10, (64,), device="cuda")
    _native.reset_launch_count()
    logp = seq(x)
    loss = F.nll_loss(logp.float(), t)
    loss.backward()
    table = _native.launch_table()
    assert table.get("gemm_bf16_tcgen05", 0) >= 6 and table.get("maxpool_fwd_nhwc", 0) == 2 and table.get("im2col_nhwc", 0) >= 1, table
    got = {n: p.grad.clone() for n, p in m.named_parameters()}
    m.zero_grad()
    ref = m.reference_logp(x, emulate_bf16=True)
    F.nll_loss(ref, t).backward()
    assert (logp.float() - ref.detach()).abs().max().item() < 0.06
    # gradients of this network at initialisation are noise-dominated (arg-max / ReLU flips on bf16 near-ties):
    # the hand-fused CifarCNN shows the same 2-10 % against the reference (tests/test_gpu_cnn.py)
    for n, p in m.named_parameters():
        assert _rel(got[n], p.grad) < (0.15 if n.startswith("conv1") else 0.08), (n, _rel(got[n], p.grad))


def test_mp_stack_fuses_itself_on_first_cuda_forward():
    import os
    import tadnn.mpi as mpi
    from tadnn import nn
    os.environ.setdefault("MASTER_PORT", "29579")
    h = mpi.TorchMPI().start(True)
    try:
        torch.manual_seed(0)
        model = (nn.Sequential().add(nn.MPInitialReshape(1024)).add(nn.MPInitialLinear(1024, 2048, mpi=h)).add(nn.MPTanh(mpi=h))
                 .add(nn.MPBaseLinear(2048, 10, mpi=h))).cuda()
        x = torch.randn(64, 1, 32, 32, device="cuda")
        y = model(x)
        from tadnn.parallel.tp import FusedMPMLP
        assert isinstance(model.__dict__.get("_tadnn_mp_fused"), FusedMPMLP)
        assert len(list(model.parameters())) == 4            # parameters shared, not duplicated
        ref = F.linear(torch.tanh(F.linear(x.view(64, 1024), model[1].weight, model[1].bias)), model[3].weight, model[3].bias)
        assert (y - ref).abs().max().item() < 3e-2 * max(1.0, ref.abs().max().item())
        y.sum().backward()
        assert model[1].weight.grad is not None and model[3].weight.grad is not None
        model2 = nn.Sequential().add(nn.MPInitialLinear(64, 32, mpi=h)).cuda()
        model2._tadnn_no_autofuse = True
        assert model2(torch.randn(4, 64, device="cuda")).shape == (4, 32) and "_tadnn_mp_fused" not in model2.__dict__
    finally:
        h.stop()
