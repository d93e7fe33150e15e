"""Collective kernels (csrc/comm.cu).

* 1 GPU: the all-reduce kernels run with all ranks emulated inside ONE cooperative launch
  (rank = blockIdx.y; the recipe forbids separate launches that wait on each other on one GPU).
* >= 2 GPUs: the real multi-process path (CUDA-IPC symmetric heaps) via torchrun, see
  tests/_dist_gpu_checks.py.
"""
import ctypes as C
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _heap_view(lib, comm, nbytes, owner):
    from tadnn.parallel.symm import _RawCuda
    p = C.c_void_p()
    rank = C.c_int()
    lib.tadnn_comm_info(comm, C.byref(rank), None, None, None)
    lib.tadnn_comm_heap(comm, rank, C.byref(p))
    return torch.as_tensor(_RawCuda(int(p.value), nbytes, owner), device="cuda")


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("algo", [1, 2])
def test_allreduce_emulated_single_launch(world, dtype, algo):
    from tadnn import _native
    lib = _native.lib()
    heap = 4 << 20
    comms = (C.c_void_p * world)()
    _native.check(lib.tadnn_comm_create_local(world, C.c_ulonglong(heap), comms), "create_local")
    try:
        for n in (5, 1000, 114838):
            es = 4 if dtype == torch.float32 else 2
            xs = [torch.randn(n, device="cuda").to(dtype) for _ in range(world)]
            views = [_heap_view(lib, C.c_void_p(comms[r]), heap, comms).view(dtype)[:n] for r in range(world)]
            for v, x in zip(views, xs):
                v.copy_(x)
            torch.cuda.synchronize()
            _native.check(lib.tadnn_allreduce_emulated(comms, world, C.c_ulonglong(0), C.c_ulonglong(n),
                                                       0 if dtype == torch.float32 else 1, C.c_float(1.0 / world),
                                                       algo, C.c_void_p(_native.stream_ptr())), "allreduce_emulated")
            torch.cuda.synchronize()
            ref = torch.stack([x.float() for x in xs]).sum(0) / world
            tol = 1e-5 if dtype == torch.float32 else 1e-2
            for v in views:
                assert (v.float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())
            for v in views[1:]:
                assert torch.equal(v, views[0])  # bit-identical on every rank
            del es
    finally:
        for r in range(world):
            lib.tadnn_comm_destroy(C.c_void_p(comms[r]))


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("algo", [1, 2])
def test_allreduce_sgd_emulated_single_launch(world, algo):
    """Fused gradient all-reduce-average + SGD update + gradient clear, all ranks in one cooperative launch."""
    from tadnn import _native
    lib = _native.lib()
    heap = 4 << 20
    comms = (C.c_void_p * world)()
    _native.check(lib.tadnn_comm_create_local(world, C.c_ulonglong(heap), comms), "create_local")
    try:
        for n in (64, 4864, 114944):
            p0 = torch.randn(n, device="cuda")
            gs = [torch.randn(n, device="cuda") for _ in range(world)]
            views = [_heap_view(lib, C.c_void_p(comms[r]), heap, comms).view(torch.float32) for r in range(world)]
            for v, g in zip(views, gs):
                v[:n].copy_(p0)
                v[n:2 * n].copy_(g)
            torch.cuda.synchronize()
            _native.check(lib.tadnn_allreduce_sgd_emulated(comms, world, C.c_ulonglong(4 * n), C.c_ulonglong(0), C.c_ulonglong(n),
                                                           C.c_float(1.0 / world), C.c_float(0.1), C.c_float(0.01), algo,
                                                           C.c_void_p(_native.stream_ptr())), "allreduce_sgd_emulated")
            torch.cuda.synchronize()
            ref = p0 - 0.1 * (torch.stack(gs).sum(0) / world + 0.01 * p0)
            for v in views:
                assert (v[:n] - ref).abs().max().item() < 1e-5
                assert float(v[n:2 * n].abs().sum()) == 0.0
                assert torch.equal(v[:n], views[0][:n])
    finally:
        for r in range(world):
            lib.tadnn_comm_destroy(C.c_void_p(comms[r]))


@pytest.mark.parametrize("world", [2, 8])
def test_allreduce_ll_emulated_single_launch(world):
    """Low-latency all-reduce ({value, epoch} pairs, no barrier), plain and fused-SGD forms, several epochs in a
    row (both parities of the double-buffered staging area), all ranks in one cooperative launch."""
    from tadnn import _native
    lib = _native.lib()
    from tadnn.parallel.symm import LL_MAX_FLOATS as LL
    ll_bytes = 2 * world * LL * 8
    heap = ll_bytes + (8 << 20)
    comms = (C.c_void_p * world)()
    _native.check(lib.tadnn_comm_create_local(world, C.c_ulonglong(heap), comms), "create_local")
    try:
        views = [_heap_view(lib, C.c_void_p(comms[r]), heap, comms) for r in range(world)]
        for v in views:
            v[:ll_bytes].zero_()
        f32 = [v[ll_bytes:].view(torch.float32) for v in views]
        epoch = 0
        for n in (1, 10, 2500, LL):
            xs = [torch.randn(n, device="cuda") for _ in range(world)]
            for v, x in zip(f32, xs):
                v[:n].copy_(x)
            torch.cuda.synchronize()
            epoch += 1
            _native.check(lib.tadnn_allreduce_ll_emulated(comms, world, C.c_ulonglong(0), C.c_ulonglong(ll_bytes), C.c_ulonglong(0),
                                                          C.c_ulonglong(n), C.c_float(1.0 / world), 0, C.c_float(0), C.c_float(0),
                                                          C.c_uint(epoch), C.c_void_p(_native.stream_ptr())), "ll")
            torch.cuda.synchronize()
            ref = torch.stack(xs).sum(0) / world
            for v in f32:
                assert (v[:n] - ref).abs().max().item() < 1e-5 and torch.equal(v[:n], f32[0][:n])
        n = 4864
        p0 = torch.randn(n, device="cuda")
        gs = [torch.randn(n, device="cuda") for _ in range(world)]
        for v, g in zip(f32, gs):
            v[:n].copy_(g)
            v[LL:LL + n].copy_(p0)
        torch.cuda.synchronize()
        epoch += 1
        _native.check(lib.tadnn_allreduce_ll_emulated(comms, world, C.c_ulonglong(0), C.c_ulonglong(ll_bytes), C.c_ulonglong(ll_bytes + 4 * LL),
                                                      C.c_ulonglong(n), C.c_float(1.0 / world), 1, C.c_float(0.1), C.c_float(0.01),
                                                      C.c_uint(epoch), C.c_void_p(_native.stream_ptr())), "ll sgd")
        torch.cuda.synchronize()
        ref = p0 - 0.1 * (torch.stack(gs).sum(0) / world + 0.01 * p0)
        for v in f32:
            assert (v[LL:LL + n] - ref).abs().max().item() < 1e-5 and float(v[:n].abs().sum()) == 0.0
    finally:
        for r in range(world):
            lib.tadnn_comm_destroy(C.c_void_p(comms[r]))


def test_world1_context_and_flat_buffer_sync():
    """parallelize() on a single GPU: flat symmetric buffer, hooks fire, sync is a no-op average."""
    import tadnn
    import tadnn.mpi as mpi
    from tadnn.models.cifar_cnn import CifarCNN
    os.environ.setdefault("MASTER_PORT", "29577")
    h = mpi.TorchMPI().start(True)
    try:
        assert h.symm is not None and h.size() == 1
        m = CifarCNN().cuda()
        x = torch.randn(32, 3, 32, 32).bfloat16()
        y = torch.randint(0, 10, (32,))
        d, t, n = tadnn.parallelize(x, y, m, 32, h, None, 2, verbose=False)
        assert n == 32 and m._tadnn_flat.symmetric
        tr = tadnn.BatchTrainer(m, 0.01)
        before = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).clone()
        for _ in range(4):
            tr.step(d.cuda(), t.cuda())
        torch.cuda.synchronize()
        assert m._tadnn_sync.syncs == 2 and m.sync_counter == 5
        after = torch.cat([p.detach().reshape(-1) for p in m.parameters()])
        assert not torch.equal(before, after)
    finally:
        h.stop()


@pytest.mark.multigpu
def test_multiprocess_collectives_via_torchrun():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29561", os.path.join(ROOT, "tests", "_dist_gpu_checks.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    line = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert p.returncode == 0 and line, p.stdout[-2000:] + p.stderr[-2000:]
    res = json.loads(line[-1])
    assert res["world"] == world and res["failures"] == []


def test_fused_mp_mlp_world1_matches_dense():
    """FusedMPMLP on one GPU (P=1): GEMM -> reduce-scatter push into the own staging slot ->
    reduce_slots(+tanh), fused tanh-backward dgrad, wgrad; vs dense fp32 autograd."""
    import tadnn.mpi as mpi
    from tadnn.parallel.tp import FusedMPMLP
    os.environ.setdefault("MASTER_PORT", "29578")
    h = mpi.TorchMPI().start(True)
    try:
        torch.manual_seed(0)
        sizes, B = [256, 512, 128, 10], 200
        m = FusedMPMLP(sizes, mpi=h).cuda()
        x = torch.randn(B, sizes[0], device="cuda")
        tgt = torch.randn(B, sizes[-1], device="cuda")
        y = m(x.bfloat16())
        ((y - tgt) ** 2).mean().backward()
        a = x.bfloat16().float()
        Ws = [w.detach().clone().requires_grad_(True) for w in m.weights]
        bs = [b.detach().clone().requires_grad_(True) for b in m.biases]
        for l in range(3):
            a = a @ Ws[l].t() + bs[l]
            if l < 2:
                a = torch.tanh(a)
        ((a - tgt) ** 2).mean().backward()
        assert (y - a.detach()).abs().max().item() < 3e-2 * max(1.0, a.abs().max().item())
        for l in range(3):
            rel = ((m.weights[l].grad - Ws[l].grad).norm() / Ws[l].grad.norm()).item()
            relb = ((m.biases[l].grad - bs[l].grad).norm() / bs[l].grad.norm()).item()
            assert rel < 4e-2 and relb < 4e-2, (l, rel, relb)
    finally:
        h.stop()
