"""``optim.ConfusionMatrix`` as used by the example
(cifar_example/sgd-torchad_nn-cifar.lua:266-275): ``add(prediction, target)``, ``totalValid``,
per-class accuracies, printable."""
from __future__ import annotations

import torch


class ConfusionMatrix:
    def __init__(self, classes):
        self.classes = list(classes) if not isinstance(classes, int) else [str(i) for i in range(classes)]
        self.nclasses = len(self.classes)
        self.mat = torch.zeros(self.nclasses, self.nclasses, dtype=torch.long)

    def zero(self):
        self.mat.zero_()

    def add(self, prediction, target):
        """``prediction``: class scores ``[C]`` / ``[B, C]`` or class indices; ``target``: 0-based."""
        p = torch.as_tensor(prediction)
        if p.is_floating_point():
            p = p.reshape(-1, self.nclasses).argmax(dim=-1)
        p = p.reshape(-1).long().cpu()
        t = torch.as_tensor(target).reshape(-1).long().cpu()
        self.mat.index_put_((t, p), torch.ones_like(t), accumulate=True)

    batchAdd = add

    @property
    def totalValid(self) -> float:
        tot = self.mat.sum().item()
        return self.mat.diag().sum().item() / tot if tot else 0.0

    @property
    def valids(self):
        return self.mat.diag().double() / self.mat.sum(1).clamp(min=1).double()

    @property
    def averageValid(self) -> float:
        return float(self.valids.mean())

    def __str__(self):
        lines = ["ConfusionMatrix:"]
        for i, row in enumerate(self.mat.tolist()):
            lines.append("[" + " ".join(f"{v:7d}" for v in row) + f"]  {100 * float(self.valids[i]):6.2f}%  [class: {self.classes[i]}]")
        lines.append(f" + average row correct: {100 * self.averageValid:.4f}%")
        lines.append(f" + global correct: {100 * self.totalValid:.4f}%")
        return "\n".join(lines)
