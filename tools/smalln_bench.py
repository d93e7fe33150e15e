"""Classifier-sized linear layer (N = 10): GEMV-style kernels (csrc/smalln.cu) vs the tcgen05 GEMM path
they replace in FusedMPMLP.  CUDA-event timed around CUDA-graph replays of 20 calls (device time per call, no launch gaps)."""
import json
import sys

import torch

from tadnn.ops import smalln
from tadnn.ops.cnn import auto_split_k, colsum_bf16
from tadnn.ops.gemm import gemm_bf16


def timed(fn, reps=20, iters=20, warm=3):
    """Per-call device time: `reps` calls captured in one CUDA graph (no host launch gaps), median over
    `iters` replays.  Falls back to back-to-back eager launches if the capture fails."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    graph = None
    try:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                for _ in range(reps):
                    fn()
        torch.cuda.synchronize()
        graph = g
    except Exception as e:  # noqa: BLE001
        print("graph capture failed, eager timing:", repr(e)[:200], flush=True)
        torch.cuda.synchronize()

    def run():
        if graph is not None:
            graph.replay()
        else:
            for _ in range(reps):
                fn()

    run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        run()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3 / reps)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    out = []
    for M, N, K in [(4096, 10, 256), (4096, 10, 2048), (32768, 10, 1024)]:
        x = torch.randn(M, K, device="cuda").bfloat16()
        w = torch.randn(N, K, device="cuda").bfloat16()
        b = torch.randn(N, device="cuda")
        gs = torch.randn(M, 16, device="cuda").bfloat16()
        gs[:, N:] = 0
        g = gs[:, :N]
        aux = torch.tanh(torch.randn(M, K, device="cuda")).bfloat16()
        y = torch.empty(M, N, device="cuda")
        db = torch.zeros(N, device="cuda")

        def wgrad_gemm():
            gemm_bf16(g, x, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32,
                      split_k=auto_split_k(N, K, M, 256), tile_n=256 if K > 128 else 128)
            db.zero_()
            colsum_bf16(g, db)

        row = {
            "M": M, "N": N, "K": K,
            "fwd_gemm_us": timed(lambda: gemm_bf16(x, w, bias=b, out=y, out_dtype=torch.float32, tile_n=64)),
            "fwd_smalln_us": timed(lambda: smalln.linear_smalln_fwd(x, w, bias=b, out=y)),
            "dgrad_gemm_us": timed(lambda: gemm_bf16(g, w, b_mn_major=True, aux_tanh_out=aux, tile_n=256 if K >= 256 else 0)),
            "dgrad_smalln_us": timed(lambda: smalln.linear_smalln_dgrad(g, w, aux_tanh_out=aux)),
            "wgrad_gemm_colsum_us": timed(wgrad_gemm),
            "wgrad_smalln_us": timed(lambda: smalln.linear_smalln_wgrad(g, x)),
        }
        row = {k: (round(v, 2) if isinstance(v, float) else v) for k, v in row.items()}
        out.append(row)
        print(json.dumps(row), flush=True)
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as fh:
            json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
