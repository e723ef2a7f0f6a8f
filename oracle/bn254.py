"""BN254 instance of oracle.curve (kept as the historical import name)."""
from .curve import *  # noqa: F401,F403
from .curve import _Fp, _Fp2  # noqa: F401
