"""Importable alias of the package directory `paritytech-bn_b200/` (a hyphen is not a valid
Python identifier).  All code lives there; this shim only extends the search path."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "paritytech-bn_b200"))

from .api import *  # noqa: E402,F401,F403
from .api import Engine, G2Precomp, BnError, CurveError, load_library, library_path, pairing_product_sharded, pairing_sharded, CtxOptions  # noqa: E402,F401
from .distributed import shard_range, combine_partials, pairing_batch_distributed  # noqa: E402,F401
