"""``nn.StochasticGradient`` with the reference's synchronisation hook built in.

The reference *redefines* ``nn.StochasticGradient:train`` class-wide (datamodule.lua:117-184):
the stock per-sample loop — ``criterion:forward(module:forward(x), y)``,
``module:updateGradInput``, ``module:accUpdateGradParameters(x, dLoss, lr)`` (an in-place SGD
step) — plus a counter that calls ``synchronizeModel`` every ``batchSize`` samples and when the
counter equals the shard size (datamodule.lua:146-158).  Here the trainer is hook-aware from
the start: ``optimize_sync`` attaches a ``SyncState`` to the model and ``train`` consults it.

Trainer fields kept: ``learningRate``, ``learningRateDecay``, ``maxIteration``,
``shuffleIndices``, ``verbose``, ``hookExample``, ``hookIteration`` (datamodule.lua:119-128,
162-182).  Extension for throughput on B200: ``miniBatch`` > 1 processes that many samples per
step (one fused forward/backward); the sync counter still counts *steps*, like the reference
counts calls.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch


class _TensorPairDataset:
    def __init__(self, data, targets):
        self.data, self.targets = data, targets

    def size(self):
        return int(self.data.shape[0])

    def __getitem__(self, i):
        return self.data[i], self.targets[i]


def _dataset_size(ds) -> int:
    if hasattr(ds, "size") and callable(ds.size):
        return int(ds.size())
    return len(ds)


class StochasticGradient:
    def __init__(self, module: torch.nn.Module, criterion):
        self.learningRate = 0.01
        self.learningRateDecay = 0.0
        self.maxIteration = 25
        self.shuffleIndices = True
        self.module = module
        self.criterion = criterion
        self.verbose = True
        self.hookExample: Optional[Callable] = None
        self.hookIteration: Optional[Callable] = None
        self.miniBatch = 1
        self.sync_counter: Optional[int] = None  # lives on the trainer (datamodule.lua:146-148)
        self.syncs = 0

    # -- one optimisation step on a (batch of) sample(s) --------------------------
    def _step(self, input, target, lr: float) -> float:
        module, criterion = self.module, self.criterion
        fast = getattr(module, "native_sample_step", None)
        if fast is not None and torch.is_tensor(input) and input.is_cuda and getattr(criterion, "is_cross_entropy", False) \
                and getattr(module, "_native_ok", lambda t: False)(input if input.dim() == 4 else input.unsqueeze(0)):
            # the whole per-sample step (forward, criterion, backward, in-place update) as one CUDA-graph replay
            return fast(input, target, lr)
        params = [p for p in module.parameters() if p.requires_grad]
        with torch.enable_grad():
            out = module(input)
            loss = criterion(out, target)
            grads = torch.autograd.grad(loss, params, allow_unused=True)
        with torch.no_grad():  # accUpdateGradParameters: W += -lr * dW, gradWeight untouched
            ps = [p for p, g in zip(params, grads) if g is not None]
            gs = [g.to(p.dtype) for p, g in zip(params, grads) if g is not None]
            if ps:
                torch._foreach_add_(ps, gs, alpha=-lr)
        return loss

    def train(self, dataset):
        if isinstance(dataset, (tuple, list)) and len(dataset) == 2 and torch.is_tensor(dataset[0]):
            dataset = _TensorPairDataset(*dataset)
        n = _dataset_size(dataset)
        iteration = 1
        lr = self.learningRate
        module = self.module
        st = getattr(module, "_tadnn_sync", None)
        idx = torch.randperm(n) if self.shuffleIndices else torch.arange(n)
        print("# StochasticGradient: training")
        mb = max(1, int(self.miniBatch))
        currentError = 0.0
        while True:
            err_acc = None
            t = 0
            while t < n:
                if mb == 1:
                    example = dataset[int(idx[t])]
                    inp, tgt = example[0], example[1]
                else:
                    sel = idx[t:t + mb]
                    if isinstance(dataset, _TensorPairDataset):
                        sel_d = sel.to(dataset.data.device)
                        inp, tgt = dataset.data[sel_d], dataset.targets[sel_d.to(dataset.targets.device)]
                    else:
                        ex = [dataset[int(i)] for i in sel]
                        inp = torch.stack([e[0] for e in ex])
                        tgt = torch.stack([torch.as_tensor(e[1]) for e in ex])
                    example = (inp, tgt)
                loss = self._step(inp, tgt, lr)
                w = inp.shape[0] if mb > 1 else 1
                err_acc = loss.detach() * w if err_acc is None else err_acc + loss.detach() * w
                # ---- additional sync functionality (datamodule.lua:146-158) ----
                if st is not None:
                    if self.sync_counter is None:
                        self.sync_counter = 1
                    if st.should_sync(self.sync_counter):
                        st.sync()
                        self.syncs += 1
                    self.sync_counter += 1
                if self.hookExample is not None:
                    self.hookExample(self, example)
                t += mb
            currentError = float(err_acc) / n
            if self.hookIteration is not None:
                self.hookIteration(self, iteration, currentError)
            if self.verbose:
                print(f"# current error = {currentError}")
            iteration += 1
            lr = self.learningRate / (1 + iteration * self.learningRateDecay)
            if self.maxIteration > 0 and iteration > self.maxIteration:
                print("# StochasticGradient: you have reached the maximum number of iterations")
                print(f"# training error = {currentError}")
                break
        return currentError
