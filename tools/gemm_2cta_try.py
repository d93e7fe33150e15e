import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tadnn.ops.gemm import gemm_bf16, gemm_bf16_2cta
torch.manual_seed(0)
for (M, N, K) in [(256, 256, 64), (256, 256, 512), (512, 768, 1024), (1000, 520, 200), (4096, 4096, 4096)]:
    a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    y = gemm_bf16_2cta(a, b, bias=bias, act="tanh", alpha=0.05)
    torch.cuda.synchronize()
    ref = torch.tanh(0.05 * (a.float() @ b.float().t()) + bias)
    print(M, N, K, "max err", (y.float() - ref).abs().max().item(), flush=True)
M = N = K = 8192
a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for name, fn in (("2cta", lambda: gemm_bf16_2cta(a, b, out=out)), ("1cta", lambda: gemm_bf16(a, b, out=out)), ("cublas", lambda: torch.matmul(a, b.t(), out=out))):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"{name} 8192^3 {ms:.3f} ms {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)
