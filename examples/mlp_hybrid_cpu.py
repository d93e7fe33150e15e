"""Hybrid data x model parallelism (not possible with the reference: its two modes share one communicator):
the README MLP built from the MP layers over a TP communicator, trained with parallelize() over a DP
communicator.  CPU / gloo, synthetic CIFAR-shaped data.

    examples/launch.sh 4 examples/mlp_hybrid_cpu.py --tp 2 --size 200 --batchSize 10
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn as automation  # noqa: E402
import tadnn.mpi as mpi  # noqa: E402
from tadnn import nn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=2, help="ranks per model-parallel group")
    ap.add_argument("--size", type=int, default=200)
    ap.add_argument("--batchSize", type=int, default=10)
    ap.add_argument("--iterations", type=int, default=1)
    ap.add_argument("--learningRate", type=float, default=1e-2)
    a = ap.parse_args()
    mpi.start(False)  # CPU / gloo
    dp, tp = mpi.mesh(a.tp)
    torch.manual_seed(1)
    g = torch.Generator().manual_seed(1)
    data = torch.randn(a.size, 1, 32, 32, generator=g)
    labels = torch.randint(0, 10, (a.size,), generator=g)
    model = (nn.Sequential().add(nn.MPInitialReshape(1024, mpi=tp)).add(nn.MPInitialLinear(1024, 2048, mpi=tp))
             .add(nn.MPTanh(mpi=tp)).add(nn.MPBaseLinear(2048, 10, mpi=tp)))
    # every DP replica of a weight shard starts from replica 0's values and is averaged with its peers
    data, labels, size = automation.parallelize(data, labels, model, a.size, dp, None, a.batchSize)
    trainer = nn.StochasticGradient(model, nn.CrossEntropyCriterion())
    trainer.learningRate, trainer.maxIteration, trainer.shuffleIndices = a.learningRate, a.iterations, False
    err = trainer.train((data, labels))
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    print(f"rank {mpi.rank()} (dp {dp.rank()}/{dp.size()}, tp {tp.rank()}/{tp.size()}): shard {size} samples, "
          f"{trainer.syncs} syncs, error {err:.4f}, |w| {flat.norm():.6f}")
    dp.stop()
    tp.stop()
    mpi.stop()


if __name__ == "__main__":
    main()
