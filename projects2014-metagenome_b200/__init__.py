"""B200-native `metagraph build` hot path (BOSS chunk construction).

The directory name carries a hyphen (it is fixed by the task); import it through the
`mtg_b200` shim at the repository root or importlib.import_module.
"""
from . import _lib, synth  # noqa: F401
from ._lib import MtgError  # noqa: F401
from .constructor import BOSSChunkConstructor, Chunk, IBOSSChunkConstructor  # noqa: F401
