"""B200-native batched rover stepping engine (the hot path of arthur-bouton/TF_ODE)."""
from . import abi  # noqa: F401
from .config import make_config, set_heightfield, synthetic_heightmap  # noqa: F401
from .engine import RoverBatch, TfodeError, glorot_policy, load_library  # noqa: F401
