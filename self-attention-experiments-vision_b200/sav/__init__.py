"""B200-native kernels behind the self-attention-experiments-vision module surface."""
from . import functional, linen, ops  # noqa: F401
from .arena import ParamArena, ParamTree  # noqa: F401
from .tape import Tape, Var, value_and_grad  # noqa: F401
from .train import load_params  # noqa: F401
