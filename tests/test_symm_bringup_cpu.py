"""Host-side control flow of the VMM / multicast heap bring-up (parallel/symm.py: _fd_exchange, _try_vmm)
with real processes over gloo and a stand-in for the native library: real file descriptors travel over
the AF_UNIX sockets, every rank must leave _try_vmm at the same point with the same verdict, and no
descriptor may leak — also when one rank fails half-way."""
import os
import socket

import pytest
import torch.multiprocessing as mp

from tadnn import mpi as tmpi
from tadnn.parallel.symm import SymmContext


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _FakeLib:
    """Mimics the entry points _try_vmm calls.  Descriptors are real (memfds), so SCM_RIGHTS passing and
    ownership are exercised; `fail` injects a failure into one phase on this rank."""

    def __init__(self, rank, world, fail=None, multicast=False):
        self.rank, self.world, self.fail, self.multicast = rank, world, fail, multicast
        self.imported, self.calls = [], []

    def _fd(self):
        fd = os.memfd_create(f"tadnn-test-{self.rank}")  # like a cuMem handle: one object, many receivers
        os.write(fd, f"rank{self.rank}".encode())
        return fd

    def tadnn_last_error(self):
        return f"injected {self.fail}".encode()

    def tadnn_comm_create_vmm(self, rank, world, heap_bytes, out):
        self.calls.append("create")
        return 1 if self.fail == "create" else 0

    def tadnn_comm_export_fd(self, handle, fd_ref):
        self.calls.append("export")
        fd_ref._obj.value = self._fd()
        return 0

    def tadnn_comm_import_peer_fd(self, handle, peer, fd):
        self.calls.append("import")
        data = os.pread(fd.value, 16, 0).decode()
        os.close(fd.value)  # the native call owns the descriptor
        self.imported.append((peer.value, data))
        return 1 if self.fail == "import" else 0

    def tadnn_multicast_supported(self, out_ref):
        out_ref._obj.value = 1 if self.multicast else 0
        return 0

    def tadnn_comm_mc_create(self, handle, fd_ref):
        self.calls.append("mc_create")
        fd_ref._obj.value = self._fd()
        return 0

    def tadnn_comm_mc_import(self, handle, fd):
        self.calls.append("mc_import")
        os.close(fd.value)
        return 0

    def tadnn_comm_mc_add_device(self, handle):
        self.calls.append("mc_add")
        return 1 if self.fail == "mc_add" else 0

    def tadnn_comm_mc_bind(self, handle):
        self.calls.append("mc_bind")
        return 0


def _open_fds():
    return set(os.listdir("/proc/self/fd"))


def _worker(rank, world, port, fail_rank, fail, multicast, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    h = tmpi.TorchMPI()
    h.start(False, backend="gloo", rank=rank, world_size=world)
    ctx = SymmContext()
    ctx.rank, ctx.world, ctx.heap_bytes = rank, world, 1 << 20
    lib = _FakeLib(rank, world, fail if rank == fail_rank else None, multicast)
    before = _open_fds()
    ok, why = ctx._try_vmm(lib)
    leaked = sorted(_open_fds() - before)
    q.put((rank, ok, why, sorted(lib.imported), lib.calls, ctx.has_multicast, leaked))
    h.stop()


def _run(world, fail_rank=None, fail=None, multicast=False):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fail_rank, fail, multicast, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    return res


@pytest.mark.parametrize("world", [2, 3])
def test_vmm_bringup_passes_descriptors_between_ranks(world):
    for rank, ok, why, imported, calls, has_mc, leaked in _run(world):
        assert ok and why == ""
        assert imported == [(q, f"rank{q}") for q in range(world) if q != rank]  # each peer's own descriptor arrived
        assert not has_mc and not leaked, (rank, leaked)


def test_vmm_bringup_with_multicast_votes():
    res = _run(2, multicast=True)
    assert all(r[1] and r[5] for r in res)
    assert "mc_create" in res[0][4] and "mc_import" in res[1][4] and all("mc_bind" in r[4] for r in res)
    assert not any(r[6] for r in res)


@pytest.mark.parametrize("fail", ["create", "import"])
def test_vmm_bringup_failure_on_one_rank_is_unanimous(fail):
    res = _run(2, fail_rank=1, fail=fail)
    assert [r[1] for r in res] == [False, False]  # both ranks fall back together
    assert "injected" in res[1][2] and res[0][2]   # the failing rank reports its error, the peer a reason
    assert not any(r[6] for r in res), [r[6] for r in res]
    if fail == "create":
        assert "import" not in res[0][4]  # nobody went on to the exchange


def test_multicast_failure_keeps_the_heap():
    res = _run(2, fail_rank=0, fail="mc_add", multicast=True)
    assert all(r[1] for r in res) and not any(r[5] for r in res)  # VMM heap stays, NVLS is off everywhere
    assert not any("mc_bind" in r[4] for r in res)
