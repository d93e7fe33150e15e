"""``optim.ConfusionMatrix`` as the reference example uses it (cifar_example/sgd-torchad_nn-cifar.lua:266-275).
Optimisers themselves are ``torch.optim`` (the models are ``torch.nn.Module``s) or the in-place SGD of
``nn.StochasticGradient`` / ``tadnn.BatchTrainer``."""
from .utils.confusion import ConfusionMatrix

__all__ = ["ConfusionMatrix"]
