"""Config 1: the README MLP (Reshape 1024 -> Linear 1024x2048 -> Tanh -> Linear 2048x10), automatic
data parallelism on CPU with gloo, synthetic CIFAR-shaped (one 32x32 plane) data.

    examples/launch.sh 2 examples/mlp_dp_cpu.py --size 200 --batchSize 10

Mirrors README.md:37-75: build the model, call parallelize(), train with StochasticGradient;
the hooked trainer synchronises the model every batchSize samples.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn as automation  # noqa: E402
import tadnn.mpi as mpi  # noqa: E402
from tadnn import nn  # noqa: E402
from tadnn.models import readme_mlp  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=200)
    ap.add_argument("--batchSize", type=int, default=None)
    ap.add_argument("--iterations", type=int, default=1)
    ap.add_argument("--learningRate", type=float, default=1e-2)
    a = ap.parse_args()
    mpi.start(False)  # CPU / gloo
    torch.manual_seed(1)
    g = torch.Generator().manual_seed(1)
    data = torch.randn(a.size, 1, 32, 32, generator=g)
    labels = torch.randint(0, 10, (a.size,), generator=g)
    model = readme_mlp()
    data, labels, size = automation.parallelize(data, labels, model, a.size, mpi.current(), None, a.batchSize)
    trainer = nn.StochasticGradient(model, nn.CrossEntropyCriterion())
    trainer.learningRate, trainer.maxIteration, trainer.shuffleIndices = a.learningRate, a.iterations, False
    err = trainer.train((data, labels))
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    print(f"rank {mpi.rank()}: shard {size} samples, {trainer.syncs} syncs, error {err:.4f}, |w| {flat.norm():.6f}")
    mpi.stop()


if __name__ == "__main__":
    main()
