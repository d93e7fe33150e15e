import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("MASTER_PORT", "29588"); os.environ.setdefault("TADNN_SYMM_HEAP_MB", "2048")
import torch
import tadnn.mpi as mpi
from tadnn.parallel.tp import FusedMPMLP
h = mpi.TorchMPI().start(True)
B = 4096
m = FusedMPMLP([2048, 16384], mpi=h).cuda()
x = torch.randn(B, 2048, device="cuda").bfloat16()
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
with torch.no_grad():
    print("forward 1 layer P=1 (cast + GEMM rs-push + reduce_slots):", round(t(lambda: m(x)), 3), "ms")
from tadnn.parallel.tp import _cast_bf16
print("cast:", round(t(lambda: _cast_bf16(m.weights[0])), 3))
from tadnn.ops.gemm import gemm_bf16
wb = _cast_bf16(m.weights[0]); out = torch.empty(B, 16384, device="cuda", dtype=torch.bfloat16)
print("gemm plain (2cta):", round(t(lambda: gemm_bf16(x, wb, bias=m.biases[0], out=out)), 3))
print("gemm 1cta tile256:", round(t(lambda: gemm_bf16(x, wb, bias=m.biases[0], out=out, tile_n=256)), 3))
slot = torch.empty(B, 16384, device="cuda", dtype=torch.bfloat16)
print("gemm rs push to local slot:", round(t(lambda: gemm_bf16(x, wb, bias=m.biases[0], push=[slot.data_ptr()], push_ldc=16384, rs_width=16384)), 3))
h.stop()
