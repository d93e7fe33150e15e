"""Literal CPU restatement of the reference's algorithms, quirks included — the test oracle.

Nothing here is used by the product path.  Each class/method mirrors one Lua function and is
written against a TorchMPI-shaped handle (``rank()``, ``size()``, ``allreduceTensor``,
``allgatherTensor``, ``barrier``), normally a ``tadnn.mpi.FakeMPI`` so that P ranks run as P
threads of one process (SURVEY §4 "Oracle").

Torch7 module protocol reproduced by hand (no autograd): ``updateOutput``,
``updateGradInput``, ``accGradParameters(input, gradOutput, scale)``,
``accUpdateGradParameters(input, gradOutput, lr)`` (= accGradParameters with scale −lr writing
into the weights themselves).

Quirks kept on purpose: Q1 sum-allreduce of gradInput slices; Q2 ``MPTanh`` /
``MPBaseReshape`` tile the slice P times; Q3 ``MPBaseLinear`` has no ``accGradParameters``
override and relies on ``MPTanh.output`` having been narrowed in place; Q6 ``size mod P``
samples dropped; Q8 counter semantics.
"""
from __future__ import annotations

import math
from typing import List

import torch


# ---------------------------------------------------------------------------- datamodule.lua
def data_parallel(data, size: int, mpi):
    """datamodule.lua:235-267 with its 1-based inclusive arithmetic."""
    remainder = 0
    stripe = (size - (size % mpi.size())) // mpi.size()
    start = (mpi.rank() * stripe) + 1
    finish = start + (stripe - 1) + remainder
    newdata = data[start - 1:finish]  # data[{{start,finish}}]
    return newdata, stripe + remainder


def heuristic(size: int) -> int:
    """datamodule.lua:68-78."""
    if size < 1000:
        return 1
    elif size < 2500:
        return 10
    elif size < 5000:
        return 50
    return 100


def synchronize_model(params: List[torch.Tensor], grads: List[torch.Tensor], mpi) -> None:
    """datamodule.lua:211-224: per-tensor allreduce then div, params first, then grads."""
    for w in params:
        mpi.allreduceTensor(w)
        w.div_(mpi.size())
    for gw in grads:
        mpi.allreduceTensor(gw)
        gw.div_(mpi.size())


def fires(counter: int, batchSize: int, size: int) -> bool:
    """datamodule.lua:102,151."""
    return (counter % batchSize == 0) or (counter == size)


# ---------------------------------------------------------------------------- Torch7 nn by hand
class OModule:
    def __init__(self):
        self.output = None
        self.gradInput = None

    def parameters(self):
        return [], []

    def accGradParameters(self, input, gradOutput, scale=1.0):
        pass

    def accUpdateGradParameters(self, input, gradOutput, lr):
        pass


class OLinear(OModule):
    def __init__(self, i, o, gen=None):
        super().__init__()
        stdv = 1.0 / math.sqrt(i)
        self.weight = (torch.rand(o, i, generator=gen) * 2 - 1) * stdv
        self.bias = (torch.rand(o, generator=gen) * 2 - 1) * stdv
        self.gradWeight = torch.zeros(o, i)
        self.gradBias = torch.zeros(o)

    def parameters(self):
        return [self.weight, self.bias], [self.gradWeight, self.gradBias]

    def _linear_out(self, x):
        return x @ self.weight.t() + self.bias

    def updateOutput(self, input):
        self.output = self._linear_out(input)
        return self.output

    def _linear_gi(self, gradOutput):
        return gradOutput @ self.weight

    def updateGradInput(self, input, gradOutput):
        self.gradInput = self._linear_gi(gradOutput)
        return self.gradInput

    def _acc(self, W, b, x, g, scale):
        if x.dim() == 1:
            W.add_(torch.outer(g, x), alpha=scale)
            b.add_(g, alpha=scale)
        else:
            W.add_(g.t() @ x, alpha=scale)
            b.add_(g.sum(0), alpha=scale)

    def accGradParameters(self, input, gradOutput, scale=1.0):
        self._acc(self.gradWeight, self.gradBias, input, gradOutput, scale)

    def accUpdateGradParameters(self, input, gradOutput, lr):
        self._acc(self.weight, self.bias, input, gradOutput, -lr)


def narrow_input(x, mpi):
    """nodemodule.lua:16-21."""
    n = x.shape[-1]
    assert n % mpi.size() == 0
    w = n // mpi.size()
    return x.narrow(x.dim() - 1, mpi.rank() * w, w)


class OMPInitialLinear(OLinear):
    """nodemodule.lua:34-73."""

    def __init__(self, i, o, mpi, gen=None):
        assert i % mpi.size() == 0
        super().__init__(i // mpi.size(), o, gen)
        self.mpi = mpi

    def updateOutput(self, input):
        x = narrow_input(input, self.mpi)
        self.output = self._linear_out(x).clone()
        self.mpi.allreduceTensor(self.output)
        return self.output

    def updateGradInput(self, input, gradOutput):
        self.gradInput = self._linear_gi(gradOutput).clone()
        self.mpi.allreduceTensor(self.gradInput)  # Q1
        return self.gradInput

    def accGradParameters(self, input, gradOutput, scale=1.0):
        OLinear.accGradParameters(self, narrow_input(input, self.mpi), gradOutput, scale)

    def accUpdateGradParameters(self, input, gradOutput, lr):
        OLinear.accUpdateGradParameters(self, narrow_input(input, self.mpi), gradOutput, lr)


class OMPBaseLinear(OLinear):
    """nodemodule.lua:86-119: same forward/updateGradInput; NO accGradParameters override (Q3):
    the caller must hand it an already-narrowed input (MPTanh.output after its
    updateGradInput)."""

    def __init__(self, i, o, mpi, gen=None):
        assert i % mpi.size() == 0
        super().__init__(i // mpi.size(), o, gen)
        self.mpi = mpi

    def updateOutput(self, input):
        x = narrow_input(input, self.mpi)
        self.output = self._linear_out(x).clone()
        self.mpi.allreduceTensor(self.output)
        return self.output

    def updateGradInput(self, input, gradOutput):
        self.gradInput = self._linear_gi(gradOutput).clone()
        self.mpi.allreduceTensor(self.gradInput)  # Q1
        return self.gradInput


class OMPTanh(OModule):
    """nodemodule.lua:132-174."""

    def __init__(self, mpi, syncTanh=False):
        super().__init__()
        self.mpi = mpi
        self.syncTanh = syncTanh

    def updateOutput(self, input):
        self.output = torch.tanh(input)
        return self.output

    def updateGradInput(self, input, gradOutput):
        self.output = narrow_input(self.output, self.mpi)  # persistent narrowing (:146)
        gi = gradOutput * (1.0 - self.output * self.output)  # THNN.Tanh_updateGradInput
        temp = gi
        for _ in range(2, self.mpi.size() + 1):
            temp = torch.cat([temp, gi], dim=-1)  # Q2 tiling
        if self.mpi.size() > 1 and self.syncTanh:
            temp = temp.clone()
            self.mpi.allgatherTensor(gi.contiguous(), temp)
        self.gradInput = temp
        return self.gradInput


class OReshape(OModule):
    def __init__(self, *size):
        super().__init__()
        self.size = tuple(size)
        self.nelement = 1
        for s in size:
            self.nelement *= s

    def updateOutput(self, input):
        if input.numel() == self.nelement and input.shape[0] != 1:
            self.output = input.reshape(self.size)
        else:
            self.output = input.reshape((input.shape[0],) + self.size)
        return self.output

    def updateGradInput(self, input, gradOutput):
        self.gradInput = gradOutput.reshape(input.shape)
        return self.gradInput


class OMPInitialReshape(OReshape):
    """nodemodule.lua:186-226: gradInput is gradOutput viewed as the NARROWED input."""

    def __init__(self, *size, mpi=None):
        super().__init__(*size)
        self.mpi = mpi

    def updateGradInput(self, input, gradOutput):
        x = narrow_input(input, self.mpi)
        self.gradInput = gradOutput.contiguous().reshape(x.shape)
        return self.gradInput


class OMPBaseReshape(OReshape):
    """nodemodule.lua:238-285: tile gradOutput P times, view as input."""

    def __init__(self, *size, mpi=None, syncReshape=False):
        super().__init__(*size)
        self.mpi = mpi
        self.syncReshape = syncReshape

    def updateGradInput(self, input, gradOutput):
        temp = gradOutput
        for _ in range(2, self.mpi.size() + 1):
            temp = torch.cat([temp, gradOutput], dim=-1)
        if self.mpi.size() > 1 and self.syncReshape:
            temp = temp.clone()
            self.mpi.allgatherTensor(gradOutput.contiguous(), temp)
        self.gradInput = temp.contiguous().reshape(input.shape)
        return self.gradInput


class OTanh(OModule):
    def updateOutput(self, input):
        self.output = torch.tanh(input)
        return self.output

    def updateGradInput(self, input, gradOutput):
        self.gradInput = gradOutput * (1.0 - self.output * self.output)
        return self.gradInput


class OSequential(OModule):
    """Torch7 ``nn.Sequential``: ``updateGradInput`` walks all modules first, then
    ``accUpdateGradParameters`` walks them again using each predecessor's *current*
    ``.output`` — the order the patched trainer relies on (datamodule.lua:141-142, Q3)."""

    def __init__(self, *mods):
        super().__init__()
        self.modules = list(mods)

    def add(self, m):
        self.modules.append(m)
        return self

    def parameters(self):
        ps, gs = [], []
        for m in self.modules:
            p, g = m.parameters()
            ps += p
            gs += g
        return ps, gs

    def updateOutput(self, input):
        cur = input
        for m in self.modules:
            cur = m.updateOutput(cur)
        self.output = cur
        return cur

    forward = updateOutput

    def updateGradInput(self, input, gradOutput):
        cur = gradOutput
        for i in range(len(self.modules) - 1, 0, -1):
            cur = self.modules[i].updateGradInput(self.modules[i - 1].output, cur)
        cur = self.modules[0].updateGradInput(input, cur)
        self.gradInput = cur
        return cur

    def accUpdateGradParameters(self, input, gradOutput, lr):
        cur = gradOutput
        for i in range(len(self.modules) - 1, 0, -1):
            self.modules[i].accUpdateGradParameters(self.modules[i - 1].output, cur, lr)
            cur = self.modules[i].gradInput
        self.modules[0].accUpdateGradParameters(input, cur, lr)

    def accGradParameters(self, input, gradOutput, scale=1.0):
        cur = gradOutput
        for i in range(len(self.modules) - 1, 0, -1):
            self.modules[i].accGradParameters(self.modules[i - 1].output, cur, scale)
            cur = self.modules[i].gradInput
        self.modules[0].accGradParameters(input, cur, scale)

    def backward(self, input, gradOutput, scale=1.0):
        """Stock ``Sequential:backward``: per module updateGradInput + accGradParameters."""
        cur = gradOutput
        for i in range(len(self.modules) - 1, 0, -1):
            prev = self.modules[i - 1].output
            g = self.modules[i].updateGradInput(prev, cur)
            self.modules[i].accGradParameters(prev, cur, scale)
            cur = g
        g = self.modules[0].updateGradInput(input, cur)
        self.modules[0].accGradParameters(input, cur, scale)
        self.gradInput = g
        return g


class OMSECriterion:
    def forward(self, input, target):
        self.output = float(((input - target) ** 2).mean())
        return self.output

    def updateGradInput(self, input, target):
        self.gradInput = 2.0 * (input - target) / input.numel()
        return self.gradInput


class OCrossEntropyCriterion:
    def forward(self, input, target):
        x = input.reshape(1, -1) if input.dim() == 1 else input
        t = torch.as_tensor(target).reshape(-1).long()
        self.output = float(torch.nn.functional.cross_entropy(x, t))
        return self.output

    def updateGradInput(self, input, target):
        x = input.reshape(1, -1) if input.dim() == 1 else input
        t = torch.as_tensor(target).reshape(-1).long()
        p = torch.softmax(x, dim=-1)
        p[torch.arange(x.shape[0]), t] -= 1.0
        self.gradInput = (p / x.shape[0]).reshape(input.shape)
        return self.gradInput


def stochastic_gradient_train(module: OSequential, criterion, data, targets, lr: float, mpi,
                              batchSize: int, size: int, max_iteration: int = 1, counter_box=None):
    """The patched ``StochasticGradient:train`` (datamodule.lua:117-184), no shuffling."""
    counter = counter_box if counter_box is not None else {"c": None}
    n = data.shape[0]
    iteration = 1
    err = 0.0
    while True:
        cur = 0.0
        for t in range(n):
            x, y = data[t], targets[t]
            cur += criterion.forward(module.forward(x), y)
            module.updateGradInput(x, criterion.updateGradInput(module.output, y))
            module.accUpdateGradParameters(x, criterion.gradInput, lr)
            if counter["c"] is None:
                counter["c"] = 1
            if fires(counter["c"], batchSize, size):
                p, g = module.parameters()
                synchronize_model(p, g, mpi)
            counter["c"] += 1
        err = cur / n
        iteration += 1
        if max_iteration > 0 and iteration > max_iteration:
            break
    return err
