"""The torch.nn (CPU) reference networks live in oracle/torch_models.py (bench.py's CPU arm times them too);
the tests keep importing them under this name."""
from oracle.torch_models import *  # noqa: F401,F403
from oracle.torch_models import Q, ident, load_into, round_bf16, set_rounding  # noqa: F401
