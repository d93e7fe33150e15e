"""Sub-communicators (mpi.split / mpi.mesh) and hybrid data x model parallelism.

The reference cannot combine parallelize() with the MP layers: both use the world communicator, so the
parameter average would mix shards of different weights (SURVEY §2.3 "Hybrid DP x TP").  With a DP and a
TP communicator they compose; the check is that a 2 x 2 hybrid run equals the 2-rank data-parallel run
of the dense statement of the same model."""
import os
import socket

import torch
import torch.multiprocessing as mp

import tadnn
from tadnn import mpi, nn

DATA = torch.randn(16, 1, 4, 4, generator=torch.Generator().manual_seed(1))
TGT = torch.randn(16, 1, 4, generator=torch.Generator().manual_seed(2))
DENSE = {  # one dense initialisation shared by every variant
    "w1": torch.rand(12, 16, generator=torch.Generator().manual_seed(3)) - 0.5,
    "b1": torch.rand(12, generator=torch.Generator().manual_seed(4)) - 0.5,
    "w2": torch.rand(4, 12, generator=torch.Generator().manual_seed(5)) - 0.5,
    "b2": torch.rand(4, generator=torch.Generator().manual_seed(6)) - 0.5,
}


class _SumBiasLinear(torch.nn.Module):
    """Dense statement of an MP linear over `tp` ranks: the full weight, and `tp` bias vectors that are
    summed — every rank of the reference holds its own full bias and all of them receive the same
    gradient (nodemodule.lua:37,90), so the effective bias moves `tp` times per step."""

    def __init__(self, w, b, tp):
        super().__init__()
        self.weight = torch.nn.Parameter(w.clone())
        self.biases = torch.nn.ParameterList([torch.nn.Parameter(b.clone() if r == 0 else torch.zeros_like(b)) for r in range(tp)])

    def forward(self, x):
        return torch.nn.functional.linear(x, self.weight, sum(self.biases))


def _dense_model(tp=2):
    return (nn.Sequential().add(nn.Reshape(16)).add(_SumBiasLinear(DENSE["w1"], DENSE["b1"], tp)).add(nn.Tanh())
            .add(_SumBiasLinear(DENSE["w2"], DENSE["b2"], tp)))


def _mp_model(tp):
    m = (nn.Sequential().add(nn.MPInitialReshape(16, mpi=tp)).add(nn.MPInitialLinear(16, 12, mpi=tp))
         .add(nn.MPTanh(mpi=tp)).add(nn.MPBaseLinear(12, 4, mpi=tp)))
    lin = [c for c in m.children() if isinstance(c, (nn.MPInitialLinear, nn.MPBaseLinear))]
    lin[0].load_dense(DENSE["w1"], DENSE["b1"], tp)
    lin[1].load_dense(DENSE["w2"], DENSE["b2"], tp)
    return m, lin


def _train(model, data, targets, handle, bs):
    d, t, n = tadnn.parallelize(data, targets, model, data.shape[0], handle, None, bs, verbose=False)
    tr = nn.StochasticGradient(model, nn.MSECriterion())
    tr.learningRate, tr.maxIteration, tr.shuffleIndices, tr.verbose = 0.05, 2, False, False
    tr.train((d, t))
    return tr.syncs


def _hybrid_body(h):
    dp, tp = mpi.mesh(2, h)
    assert (dp.size(), tp.size()) == (2, 2)
    assert tp.rank() == h.rank() % 2 and dp.rank() == h.rank() // 2
    assert tp.global_rank(1) == (h.rank() // 2) * 2 + 1 and dp.global_rank(1) == 2 + h.rank() % 2
    # collectives stay inside their communicator
    t = torch.tensor([float(h.rank())])
    tp.allreduceTensor(t)
    assert t.item() == {0: 1.0, 1: 1.0, 2: 5.0, 3: 5.0}[h.rank()]
    t = torch.tensor([float(h.rank())])
    dp.broadcastTensor(1, t)  # root is a rank of the DP communicator
    assert t.item() == 2.0 + h.rank() % 2
    out = torch.empty(2)
    dp.allgatherTensor(torch.tensor([float(h.rank())]), out)
    assert out.tolist() == [float(h.rank() % 2), float(2 + h.rank() % 2)]
    model, lin = _mp_model(tp)
    syncs = _train(model, DATA, TGT, dp, 2)
    w1, b1 = lin[0].dense_weight_bias(tp)
    w2, b2 = lin[1].dense_weight_bias(tp)
    return syncs, w1, b1, w2, b2


def _dp_body(h):
    model = _dense_model()
    syncs = _train(model, DATA, TGT, h, 2)
    lin = [c for c in model.children() if isinstance(c, _SumBiasLinear)]
    return (syncs, lin[0].weight.detach().clone(), sum(b.detach() for b in lin[0].biases),
            lin[1].weight.detach().clone(), sum(b.detach() for b in lin[1].biases))


def _check_hybrid_equals_dp(hybrid, ref):
    for r, res in enumerate(hybrid):
        want = ref[r // 2]  # DP replica r//2 saw the same shard of the data
        assert res[0] == want[0]
        for got, exp in zip(res[1:], want[1:]):
            assert torch.allclose(got, exp, atol=1e-5), (r, (got - exp).abs().max())


def test_split_and_hybrid_training_on_fake_world():
    hybrid = mpi.FakeWorld(4).run(_hybrid_body)
    ref = mpi.FakeWorld(2).run(_dp_body)
    _check_hybrid_equals_dp(hybrid, ref)


def test_split_keys_reorder_and_solo_groups():
    def body(h):
        rev = h.split(0, key=-h.rank())  # one group, reversed order
        assert rev.size() == 3 and rev.rank() == 2 - h.rank()
        solo = h.split(h.rank())  # every rank alone
        assert solo.size() == 1 and solo.rank() == 0
        t = torch.tensor([3.0])
        solo.allreduceTensor(t)
        solo.barrier()
        return t.item()

    assert mpi.FakeWorld(3).run(body) == [3.0, 3.0, 3.0]


# ---------------------------------------------------------------- real processes (gloo)
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    h = mpi.TorchMPI()
    h.start(False, backend="gloo", rank=rank, world_size=world)
    mpi.set_current(h)
    res = _hybrid_body(h) if world == 4 else _dp_body(h)
    torch.save(res, os.path.join(out_dir, f"w{world}_r{rank}.pt"))
    h.stop()


def test_hybrid_training_over_gloo_process_groups(tmp_path):
    for world in (4, 2):
        mp.spawn(_gloo_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    hybrid = [torch.load(tmp_path / f"w4_r{r}.pt") for r in range(4)]
    ref = [torch.load(tmp_path / f"w2_r{r}.pt") for r in range(2)]
    _check_hybrid_equals_dp(hybrid, ref)
