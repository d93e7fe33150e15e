"""The reference's own workload: ONE sample per step (datamodule.lua:134-158, `-batchSize 1` in
cifar_example/train.sh:14; nodemodule.lua:40-54 at B = 1).  Latency-bound: every number here is samples/s of a
per-sample training loop, device-timed, max over ranks.

    python examples/latency_bench.py                                   # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/latency_bench.py

CNN (data-parallel, model synchronised every `--sync-every` samples like the hooked trainer):
  graph     CifarCNN.native_sample_step: one CUDA-graph replay of the fused tadnn kernels + in-place update
  layers    the same network as a Sequential of stock layers (native per-layer kernels, autograd)
  pytorch   plain PyTorch + cuDNN + NCCL restatement (baseline/torch_nccl_baseline.py), per-sample
README MP-MLP (1024 -> 2048 -> 10, model-parallel over all ranks), B = 1 and 16:
  fused     FusedMPMLP (GEMM epilogue push, low-latency pair all-reduce for the classifier)
  nccl      MPInitialLinear/MPTanh/MPBaseLinear on torch.distributed
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, n, warm, P, dev):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    if P > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for i in range(n):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if P > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=400)
    ap.add_argument("--sync-every", type=int, default=10, help="the reference's heuristic for shards < 2500 samples")
    ap.add_argument("--out", default="gpurun_out")
    a = ap.parse_args()
    import tadnn
    import tadnn.mpi as mpi
    from tadnn import nn
    from tadnn.models.cifar_cnn import CifarCNN
    h = mpi.start(True)
    P, rank = h.size(), h.rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(1)
    N = 512
    res = {"world": P, "samples_timed": a.samples, "sync_every": a.sync_every}

    # ------------------------------------------------------------------ CNN, data-parallel, per sample
    x = torch.randn(N * P, 3, 32, 32).bfloat16()
    y = torch.randint(0, 10, (N * P,))
    for name in ("graph", "layers"):
        model = CifarCNN().to(dev)
        d, t, n = tadnn.parallelize(x, y, model, N * P, h, None, a.sync_every, verbose=False)
        d, t = d.to(dev), t.to(dev)
        st = getattr(model, "_tadnn_sync", None)
        if name == "graph":
            tr = nn.StochasticGradient(model, nn.CrossEntropyCriterion())
        else:
            seq = model.to_sequential()
            tr = nn.StochasticGradient(seq, nn.CrossEntropyCriterion())
            d = d.float()
        cnt = [1]

        def step(i, tr=tr, d=d, t=t, st=st, cnt=cnt):
            tr._step(d[i % n], t[i % n], 1e-3)
            if st is not None and P > 1:
                if st.should_sync(cnt[0]):
                    st.sync()
                cnt[0] += 1
        ms = timed(step, a.samples, 30, P, dev)
        res["cnn_" + name] = {"samples_per_s": a.samples * P / ms * 1e3, "us_per_sample": ms / a.samples * 1e3}
        if st is not None:
            st.remove()
    from baseline.torch_nccl_baseline import BaselineTrainer
    bt = BaselineTrainer(dev, world=1, lr=1e-3)
    bt.world = P
    xs, ys = x[rank * N:(rank + 1) * N].to(dev).float(), y[rank * N:(rank + 1) * N].to(dev)
    cnt = [1]

    def bstep(i):
        for p in bt.params:
            p.grad = None
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logp = bt.model(xs[i % N:i % N + 1])
        torch.nn.functional.nll_loss(logp, ys[i % N:i % N + 1]).backward()
        with torch.no_grad():
            torch._foreach_add_(bt.params, [p.grad for p in bt.params], alpha=-1e-3)
        if P > 1:
            if cnt[0] % a.sync_every == 0:
                bt._sync()
            cnt[0] += 1
    ms = timed(bstep, a.samples, 30, P, dev)
    res["cnn_pytorch"] = {"samples_per_s": a.samples * P / ms * 1e3, "us_per_sample": ms / a.samples * 1e3}

    # ------------------------------------------------------------------ README MP-MLP at B = 1 and 16
    if P > 1:
        from tadnn.parallel.tp import FusedMPMLP
        sizes = [1024, 2048, 10]
        for B in (1, 16):
            xb = torch.randn(64, B, sizes[0], device=dev).bfloat16()
            tg = torch.randn(B, sizes[-1], device=dev)
            for impl in ("fused", "nccl"):
                torch.manual_seed(1)
                if impl == "fused":
                    model = FusedMPMLP(sizes, mpi=h).to(dev)
                else:
                    hn = h.nccl_only()
                    model = nn.Sequential().add(nn.MPInitialLinear(sizes[0], sizes[1], mpi=hn)).add(nn.MPTanh(mpi=hn)) \
                        .add(nn.MPBaseLinear(sizes[1], sizes[2], mpi=hn)).to(dev)
                    model._tadnn_no_autofuse = True
                params = list(model.parameters())

                def mstep(i, model=model, params=params, impl=impl):
                    yv = model(xb[i % 64]) if impl == "fused" else torch.nn.Sequential.forward(model, xb[i % 64])
                    ((yv.float() - tg) ** 2).mean().backward()
                    if impl == "fused":
                        model.sgd_step(1e-3)
                    else:
                        with torch.no_grad():
                            torch._foreach_add_(params, [p.grad for p in params], alpha=-1e-3)
                            for p in params:
                                p.grad = None
                try:
                    ms = timed(mstep, a.samples // 2, 20, P, dev)
                    res[f"mp_readme_B{B}_{impl}"] = {"samples_per_s": (a.samples // 2) * B / ms * 1e3, "us_per_step": ms / (a.samples // 2) * 1e3}
                except Exception as e:  # noqa: BLE001 - keep the other rows
                    res[f"mp_readme_B{B}_{impl}"] = {"error": f"{type(e).__name__}: {e}"[:300]}
                del model, params
    if rank == 0:
        print(json.dumps(res), flush=True)
        os.makedirs(a.out, exist_ok=True)
        with open(os.path.join(a.out, f"latency_P{P}.json"), "w") as f:
            json.dump(res, f, indent=1)
    h.stop()


if __name__ == "__main__":
    main()
