"""Config 1: README MLP (Reshape 1024 -> Linear 1024x2048 -> Tanh -> Linear 2048x10), data-parallel,
world_size=2, CPU/gloo, real processes (torch.multiprocessing.spawn + TCP store on 127.0.0.1).
Final weights must equal the single-process FakeWorld run of the same script."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

import tadnn
from tadnn import nn
from tadnn import mpi as tmpi


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _build(seed):
    g = torch.Generator().manual_seed(seed)  # explicit generator: FakeWorld ranks are threads
    m = nn.Sequential().add(nn.Reshape(1024)).add(nn.Linear(1024, 2048)).add(nn.Tanh()).add(nn.Linear(2048, 10))
    with torch.no_grad():
        for p in m.parameters():
            p.copy_((torch.rand(p.shape, generator=g) - 0.5) * 0.05)
    return m


def _data():
    g = torch.Generator().manual_seed(42)
    return torch.randn(12, 1, 32, 32, generator=g), torch.randint(0, 10, (12,), generator=g)


def _train(h):
    data, labels = _data()
    model = _build(seed=10 + h.rank())
    d, t, n = tadnn.parallelize(data, labels, model, 12, h, None, 2, verbose=False)
    tr = nn.StochasticGradient(model, nn.CrossEntropyCriterion())
    tr.learningRate, tr.maxIteration, tr.shuffleIndices, tr.verbose = 0.01, 1, False, False
    tr.train((d, t))
    return torch.cat([p.detach().reshape(-1) for p in model.parameters()]), n, tr.syncs


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    h = tmpi.TorchMPI()
    h.start(False, backend="gloo", rank=rank, world_size=world)
    tmpi.set_current(h)
    flat, n, syncs = _train(h)
    # async half of the selector table (TorchMPI's handle API): start, overlap host work, then syncHandle
    t = torch.full((1000,), float(rank + 1))
    b = torch.full((7,), float(rank))
    ha = h.collectiveSelector["cpu"]["singlenode"]["async"]["allreduceTensor"](t)
    hb = h.collectiveSelector["cpu"]["singlenode"]["async"]["broadcastTensor"](1, b)
    busy = sum(i for i in range(1000))  # host work while the collectives are in flight
    h.syncHandle(ha)
    h.syncHandle(hb)
    async_ok = bool(busy and torch.all(t == sum(range(1, world + 1))) and torch.all(b == 1.0)
                    and type(ha).__name__ == "_WorkHandle")
    torch.save({"flat": flat, "n": n, "syncs": syncs, "async_ok": async_ok, "sel": callable(
        tadnn.selectCollective(flat, "sync", "allreduceTensor", h)), "inter": h.needInterNodeCollectives},
        os.path.join(out_dir, f"r{rank}.pt"))
    h.stop()


@pytest.mark.timeout(300)
def test_mlp_dp_gloo_world2_matches_fakeworld(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(tmp_path / "r0.pt")
    r1 = torch.load(tmp_path / "r1.pt")
    assert r0["n"] == r1["n"] == 6 and r0["syncs"] == r1["syncs"] == 3
    assert r0["sel"] and r0["inter"] is False
    assert r0["async_ok"] and r1["async_ok"]
    assert torch.allclose(r0["flat"], r1["flat"], atol=1e-6)  # last sample synced (counter == size)
    fake = tmpi.FakeWorld(2).run(_train)
    assert torch.allclose(fake[0][0], r0["flat"], atol=1e-5)
