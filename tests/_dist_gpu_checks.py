"""Multi-GPU numerics worker (launched by torchrun from test_gpu_multigpu.py or by hand):

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/_dist_gpu_checks.py

Checks every collective kernel of csrc/comm.cu against a plain PyTorch fp32 reference
computed from NCCL all_gather'ed inputs, then prints one JSON line from rank 0.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import tadnn.mpi as mpi  # noqa: E402
from tadnn.parallel import symm as S  # noqa: E402


def gather_all(t):
    parts = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, t.contiguous())
    return parts


def main():
    h = mpi.start(True)
    ctx = h.symm
    assert ctx is not None, "symmetric context not created"
    rank, P = h.rank(), h.size()
    dev = ctx.device
    fails = []
    checks = 0
    torch.manual_seed(1234 + rank)

    def expect(name, got, ref, tol):
        nonlocal checks
        checks += 1
        err = (got.float() - ref.float()).abs().max().item()
        lim = tol * max(1.0, ref.float().abs().max().item())
        if not (err <= lim):
            fails.append(f"{name}: err {err:.3e} > {lim:.3e}")

    # ---- all-reduce, in heap, every algo/dtype/size (incl. odd sizes and tails)
    big = ctx.empty((1 << 24,), torch.float32)  # 64 MB region reused for all sizes
    for dtype in (torch.float32, torch.bfloat16):
        for n in (1, 3, 10, 64, 1000, 4099, 65536, 114838 * 2, 1 << 20, (1 << 22) + 5):
            for algo in ((1, 2, 3) if ctx.has_multicast else (1, 2)):
                if algo == 1 and n > (1 << 20):
                    continue
                view = big.view(dtype)[:n]
                x = torch.randn(n, device=dev).to(dtype)
                view.copy_(x)
                parts = gather_all(x)
                ref = torch.stack([p.float() for p in parts]).sum(0) / P
                ctx.allreduce_(view, scale=1.0 / P, algo=algo)
                torch.cuda.synchronize()
                expect(f"allreduce n={n} {dtype} algo={algo}", view, ref, 1e-5 if dtype == torch.float32 else 8e-3)
                # bit-identical across ranks
                got_all = gather_all(view.clone())
                if any(not torch.equal(got_all[0], g) for g in got_all):
                    fails.append(f"allreduce n={n} {dtype} algo={algo}: ranks disagree bitwise")
    # low-latency path (algo 4 = {value, epoch} pairs, no barrier): many calls in a row, varying sizes
    for it in range(60):
        n = (1, 7, 640, 4864, 16384, 115000)[it % 6]
        view = big[:n]
        x = torch.randn(n, device=dev)
        view.copy_(x)
        ref = torch.stack(gather_all(x)).sum(0) / P
        ctx.allreduce_(view, scale=1.0 / P, algo=4)
        if it % 7 == 0:
            torch.cuda.synchronize()
        expect(f"allreduce LL n={n} it={it}", view, ref, 1e-5)
    # repeated launches back-to-back (flag reuse without host sync)
    view = big[:50000]
    x = torch.randn(50000, device=dev)
    view.copy_(x)
    for _ in range(200):
        ctx.allreduce_(view, scale=1.0 / P, algo=1)
    for _ in range(200):
        ctx.allreduce_(view, scale=1.0 / P, algo=2)
    torch.cuda.synchronize()
    parts = gather_all(x)
    expect("allreduce x400 (mean is idempotent after the first)", view, torch.stack(parts).sum(0) / P, 1e-4)

    # ---- staged path (tensor outside the heap) through the TorchMPI-shaped API
    y = torch.randn(777, 33, device=dev)
    ref = torch.stack(gather_all(y)).sum(0)
    h.allreduceTensor(y)
    torch.cuda.synchronize()
    expect("mpi.allreduceTensor staged", y, ref, 1e-5)

    # ---- broadcast
    reg = ctx.empty((100003,), torch.float32)
    reg.copy_(torch.full((100003,), float(rank + 1), device=dev))
    off = ctx.offset_of(reg)
    for root in range(P):
        reg.fill_(float(rank + 1))
        torch.cuda.synchronize()
        dist.barrier()
        ctx.broadcast_offset(off, reg.numel() * 4, root=root)
        torch.cuda.synchronize()
        expect(f"broadcast root={root}", reg, torch.full_like(reg, float(root + 1)), 0.0)

    # ---- all-gather along last dim
    for (rows, w, dtype) in ((1, 256, torch.float32), (37, 10, torch.float32), (128, 2048, torch.bfloat16), (5, 3, torch.bfloat16)):
        inp = torch.randn(rows, w, device=dev).to(dtype)
        out = torch.empty(rows, w * P, device=dev, dtype=dtype)
        ref = torch.cat(gather_all(inp), dim=-1)
        h.allgatherTensor(inp, out)
        torch.cuda.synchronize()
        expect(f"allgather {rows}x{w} {dtype}", out, ref, 0.0)

    # ---- reduce-scatter along last dim with fused bias + tanh
    for (rows, w, dtype) in ((64, 256, torch.float32), (33, 40, torch.bfloat16)):
        full = torch.randn(rows, w * P, device=dev).to(dtype)
        reg2 = ctx.empty((rows, w * P), dtype)
        reg2.copy_(full)
        bias = torch.randn(w * P, device=dev)
        dst = torch.empty(rows, w, device=dev, dtype=dtype)
        allp = gather_all(full)
        ref = torch.tanh(torch.stack([p.float() for p in allp]).sum(0)[:, rank * w:(rank + 1) * w] * 0.5 + bias[rank * w:(rank + 1) * w])
        torch.cuda.synchronize()
        dist.barrier()
        ctx.reduce_scatter_cols_offset(ctx.offset_of(reg2), dst, scale=0.5, bias=bias, act=1)
        torch.cuda.synchronize()
        expect(f"reduce_scatter {rows}x{w} {dtype}", dst, ref, 1e-5 if dtype == torch.float32 else 1e-2)

    # ---- fused model-parallel MLP (GEMM epilogue -> peer staging slots -> reduce) vs dense fp32
    from tadnn.parallel.tp import FusedMPMLP
    cases = [([1024, 2048, 10], 96, "push", 1), ([512, 1024, 1024, 512], 200, "push", 1)]
    if ctx.has_multicast:  # in-switch (multimem.ld_reduce) form of the same reduction, unchunked and row-pipelined
        cases += [([512, 1024, 1024, 512], 200, "nvls", 1), ([1024, 2048, 1024, 10], 1536, "nvls", 3)]
    for sizes, B, mode, chunks in cases:
        g = torch.Generator(device="cpu").manual_seed(77)
        mlp = FusedMPMLP(sizes, mpi=h).to(dev)
        mlp.reduce_mode, mlp.nvls_chunks = mode, chunks
        tag = f"fused MP-MLP[{mode}x{chunks}] {sizes}"
        dense_W, dense_b = [], []
        for l in range(len(sizes) - 1):
            Wd = (torch.rand(sizes[l + 1], sizes[l], generator=g) - 0.5) * (2.0 / sizes[l] ** 0.5)
            bd = (torch.rand(sizes[l + 1], generator=g) - 0.5) * 0.2
            sh = sizes[l] // P
            with torch.no_grad():
                mlp.weights[l].copy_(Wd[:, rank * sh:(rank + 1) * sh])
                mlp.biases[l].copy_(bd / P)  # effective bias = sum over ranks (Q4)
            dense_W.append(Wd.to(dev).requires_grad_(True))
            dense_b.append(bd.to(dev).requires_grad_(True))
        x = torch.randn(B, sizes[0], generator=g).to(dev)
        tgt = torch.randn(B, sizes[-1], generator=g).to(dev)
        y = mlp(x.bfloat16())
        loss = ((y - tgt) ** 2).mean()
        loss.backward()
        a = x.bfloat16().float()
        for l in range(len(sizes) - 1):
            a = a @ dense_W[l].t() + dense_b[l]
            if l < len(sizes) - 2:
                a = torch.tanh(a)
        ref_loss = ((a - tgt) ** 2).mean()
        ref_loss.backward()
        torch.cuda.synchronize()
        expect(f"{tag} forward", y, a.detach(), 3e-2)
        for l in range(len(sizes) - 1):
            sh = sizes[l] // P
            gw_ref = dense_W[l].grad[:, rank * sh:(rank + 1) * sh]
            rel = ((mlp.weights[l].grad - gw_ref).norm() / (gw_ref.norm() + 1e-12)).item()
            checks += 1
            if not rel < 4e-2:
                fails.append(f"{tag} layer {l} dW rel err {rel:.3e}")
            relb = ((mlp.biases[l].grad - dense_b[l].grad).norm() / (dense_b[l].grad.norm() + 1e-12)).item()
            checks += 1
            if not relb < 4e-2:
                fails.append(f"{tag} layer {l} db rel err {relb:.3e}")

    # ---- gradient all-reduce fused with the SGD step: every algo, vs the two-pass statement
    for n in (64, 4800, 114944, (1 << 20) + 64):
        for algo in ((1, 2, 3, 4) if ctx.has_multicast else (1, 2, 4)):
            if algo == 4 and n > 131072:
                continue
            if algo == 1 and n > (1 << 18):
                continue
            reg3 = big[:2 * n]
            p0 = torch.randn(n, generator=torch.Generator().manual_seed(5)).to(dev)  # identical on every rank
            g0 = torch.randn(n, device=dev)
            reg3[:n].copy_(p0)
            reg3[n:].copy_(g0)
            gavg = torch.stack(gather_all(g0)).sum(0) / P
            ref = p0 - 0.1 * (gavg + 0.01 * p0)
            torch.cuda.synchronize()
            dist.barrier()
            off = ctx.offset_of(reg3)
            ctx.allreduce_sgd_offset(off + 4 * n, off, n, 1.0 / P, 0.1, 0.01, algo=algo, channel=S.CH_DEFAULT)
            torch.cuda.synchronize()
            expect(f"allreduce_sgd n={n} algo={algo} params", reg3[:n], ref, 1e-5)
            expect(f"allreduce_sgd n={n} algo={algo} grads cleared", reg3[n:], torch.zeros(n, device=dev), 0.0)
            got_all = gather_all(reg3[:n].clone())
            if any(not torch.equal(got_all[0], g) for g in got_all):
                fails.append(f"allreduce_sgd n={n} algo={algo}: parameters differ bitwise across ranks")
            dist.barrier()

    # ---- data-parallel CNN: fused sync+SGD step (tail overlapped with conv1's backward) == unfused two-pass step
    import tadnn
    from tadnn.models.cifar_cnn import CifarCNN
    results = {}
    for fused in ("1", "0"):
        os.environ["TADNN_DP_FUSED_SGD"] = fused
        torch.manual_seed(3)
        m = CifarCNN(dropout=0.0).to(dev)
        xs = torch.randn(P * 64, 3, 32, 32, generator=torch.Generator().manual_seed(11)).bfloat16()
        ys = torch.randint(0, 10, (P * 64,), generator=torch.Generator().manual_seed(12))
        d, t, n_sh = tadnn.parallelize(xs, ys, m, P * 64, h, None, 1, verbose=False)
        tr = tadnn.BatchTrainer(m, learningRate=0.05, weightDecay=0.01)
        for _ in range(3):
            tr.step(d.to(dev), t.to(dev))
        torch.cuda.synchronize()
        results[fused] = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).clone()
        checks += 1
        if m._tadnn_sync.fused_step_ok() != (fused == "1" and (P & (P - 1)) == 0):
            fails.append(f"fused_step_ok() wrong for TADNN_DP_FUSED_SGD={fused}")
        got_all = gather_all(results[fused])
        if any(not torch.equal(got_all[0], g) for g in got_all):
            fails.append(f"DP CNN fused={fused}: replicas differ bitwise after 3 steps")
        m._tadnn_sync.remove()
    os.environ.pop("TADNN_DP_FUSED_SGD", None)
    expect("DP CNN fused vs unfused sync+SGD", results["1"], results["0"], 1e-5)

    # ---- overlap=True vs overlap=False with DIVERGED replicas ('model' mode, sync every 2nd backward)
    def run_overlap(ov):
        torch.manual_seed(9)
        net = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.Tanh(), torch.nn.Linear(512, 256), torch.nn.Tanh(),
                                  torch.nn.Linear(256, 10)).to(dev)
        xs = torch.randn(P * 32, 256, generator=torch.Generator().manual_seed(21))
        ys = torch.randint(0, 10, (P * 32,), generator=torch.Generator().manual_seed(22))
        d, t, _n = tadnn.parallelize(xs, ys, net, P * 32, h, None, 2, verbose=False, overlap=ov)
        d, t = d.to(dev), t.to(dev)
        for i in range(4):
            loss = torch.nn.functional.cross_entropy(net(d[i * 8:(i + 1) * 8]), t[i * 8:(i + 1) * 8])
            loss.backward()
            with torch.no_grad():
                for p_ in net.parameters():
                    p_.add_(p_.grad, alpha=-0.05)   # per-rank SGD on different data: replicas diverge between syncs
                    p_.grad.zero_()
        torch.cuda.synchronize()
        out = torch.cat([p_.detach().reshape(-1) for p_ in net.parameters()]).clone()
        net._tadnn_sync.remove()
        return out
    try:
        a_, b_ = run_overlap(True), run_overlap(False)
        expect("DP overlap=True vs overlap=False (diverged replicas)", a_, b_, 1e-6)
    except TypeError as e:  # parallelize() without an overlap keyword
        fails.append(f"overlap check could not run: {e}")

    # ---- device barrier stress
    for _ in range(500):
        ctx.barrier()
    torch.cuda.synchronize()

    allf = [None] * P
    dist.all_gather_object(allf, fails)
    if rank == 0:
        flat = [f"r{r}: {m}" for r, fs in enumerate(allf) for m in fs]
        print(json.dumps({"world": P, "checks": checks, "failures": flat, "backend": ctx.backend, "multicast": ctx.has_multicast}))
    h.stop()
    sys.exit(1 if any(allf) else 0)


if __name__ == "__main__":
    main()
