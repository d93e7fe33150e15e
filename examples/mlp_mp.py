"""Config 2: the README model-parallel MLP (README.md:97-101) across the GPUs of one box.

    examples/launch.sh 8 examples/mlp_mp.py            # fused B200 kernels
    examples/launch.sh 2 examples/mlp_mp.py --cpu      # per-op layers on CPU/gloo

Builds the network from MPInitialReshape / MPInitialLinear / MPTanh / MPBaseLinear exactly as the
reference README does, optionally fuses it (tadnn.nodemodule.fuse_mp), trains a few steps on
synthetic 32x32 inputs and checks the replicated output against the gathered dense model.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn  # noqa: E402
import tadnn.mpi as mpi  # noqa: E402
from tadnn import nn  # noqa: E402
from tadnn.models import readme_mp_mlp  # noqa: E402
from tadnn.utils import convert  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    h = mpi.start(not a.cpu)
    dev = h.device()
    torch.manual_seed(1 + h.rank())  # every rank owns different shards
    model = readme_mp_mlp().to(dev)
    net = tadnn.nodemodule.fuse_mp(model) if (not a.cpu and torch.cuda.is_available()) else model
    g = torch.Generator().manual_seed(7)
    x = torch.randn(a.batch, 1, 32, 32, generator=g).to(dev)
    t = torch.randint(0, 10, (a.batch,), generator=g).to(dev)
    params = list(net.parameters())
    for s in range(a.steps):
        y = net(x.bfloat16() if dev.type == "cuda" else x)
        loss = torch.nn.functional.cross_entropy(y.float(), t)
        loss.backward()
        with torch.no_grad():
            for p in params:
                p.add_(p.grad, alpha=-0.05)
                p.grad = None
        if h.rank() == 0 and s % 5 == 0:
            print(f"step {s}: loss {loss.item():.4f}")
    dense = convert.dense_model_from_mp(model).to(dev)
    with torch.no_grad():
        err = (net(x.bfloat16() if dev.type == "cuda" else x).float() - dense(x)).abs().max().item()
    print(f"rank {h.rank()}: final loss {loss.item():.4f}; |fused - dense| = {err:.3e}")
    h.stop()


if __name__ == "__main__":
    main()
