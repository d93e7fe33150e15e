"""Fused MSE criterion (csrc/loss.cu) vs plain PyTorch fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,tdtype", [((64, 128), torch.float32), ((4096, 1024), torch.bfloat16), ((7, 13), torch.float32)])
def test_mse_fused_matches_torch(shape, tdtype):
    from tadnn.ops.loss import mse_fused
    torch.manual_seed(0)
    y = torch.randn(shape, device="cuda").bfloat16().requires_grad_(True)
    t = torch.randn(shape, device="cuda").to(tdtype)
    loss = mse_fused(y, t)
    (3.0 * loss).backward()
    yr = y.detach().float().requires_grad_(True)
    ref = torch.nn.functional.mse_loss(yr, t.float())
    (3.0 * ref).backward()
    assert abs(loss.item() - ref.item()) < 1e-4 * max(1.0, ref.item())
    assert (y.grad.float() - yr.grad).abs().max().item() < 1e-2 * yr.grad.abs().max().item() + 1e-8
    loss2 = mse_fused(y.detach(), t)  # second call in a row: the loss cell is re-zeroed
    assert abs(loss2.item() - ref.item()) < 1e-4 * max(1.0, ref.item())
