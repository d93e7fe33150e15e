"""Model zoo: the networks the reference and BASELINE.json name."""
from .cifar_cnn import CifarCNN  # noqa: F401
from .mlp import readme_mlp, readme_mp_mlp, wide_mp_mlp  # noqa: F401
