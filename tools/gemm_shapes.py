import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tadnn.ops.gemm import gemm_bf16
from tadnn.ops.cnn import auto_split_k, colsum_bf16
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
B, O, I = 4096, 16384, 2048
x = torch.randn(B, I, device="cuda").bfloat16(); W = torch.randn(O, I, device="cuda").bfloat16()
bias = torch.randn(O, device="cuda")
g = torch.randn(B, O, device="cuda").bfloat16(); y = torch.randn(B, I, device="cuda").bfloat16()
out = torch.empty(B, O, device="cuda", dtype=torch.bfloat16)
fl = 2.0 * B * O * I
ms = t(lambda: gemm_bf16(x, W, bias=bias, out=out)); print(f"fwd  [4096x2048]x[16384x2048]^T  {ms:.3f} ms {fl/ms/1e9:.0f} TF")
ms = t(lambda: torch.matmul(x, W.t())); print(f"fwd  cuBLAS                       {ms:.3f} ms {fl/ms/1e9:.0f} TF")
for tn in (128, 256):
    ms = t(lambda: gemm_bf16(g, W, b_mn_major=True, aux_tanh_out=y, tile_n=tn)); print(f"dgrad tile_n={tn}                 {ms:.3f} ms {fl/ms/1e9:.0f} TF")
ms = t(lambda: torch.matmul(g, W)); print(f"dgrad cuBLAS                      {ms:.3f} ms {fl/ms/1e9:.0f} TF")
ms = t(lambda: gemm_bf16(g, x, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32, split_k=auto_split_k(O, I, B, 256))); print(f"wgrad fp32 out                    {ms:.3f} ms {fl/ms/1e9:.0f} TF")
ms = t(lambda: torch.matmul(g.t(), x)); print(f"wgrad cuBLAS (bf16 out)           {ms:.3f} ms {fl/ms/1e9:.0f} TF")
db = torch.zeros(O, device="cuda")
ms = t(lambda: colsum_bf16(g, db)); print(f"colsum [4096,16384]               {ms:.3f} ms")
Wf = torch.randn(O, I, device="cuda")
from tadnn.parallel.tp import _cast_bf16
ms = t(lambda: _cast_bf16(Wf)); print(f"cast fp32->bf16 [16384,2048]      {ms:.3f} ms")
