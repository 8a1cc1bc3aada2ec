"""B200-native batched rover stepping engine (the hot path of arthur-bouton/TF_ODE)."""
from . import abi  # noqa: F401
from .config import PRESET_MT_TREE, make_config, reference_heightmap, set_heightfield, synthetic_heightmap  # noqa: F401
from .engine import RoverBatch, TfodeError, format_transitions, glorot_policy, load_heightmap_png, load_library, parse_model_tree, read_saved_model_actor  # noqa: F401
