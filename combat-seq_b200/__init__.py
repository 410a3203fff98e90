"""combat-seq_b200 — B200-native ComBat-seq engine (hand-written sm_100a kernels behind a C ABI).

The directory name carries a hyphen, so load it with `__graft_entry__.load_package()` (importlib) or add the
repo root to sys.path and use `importlib.import_module("combat-seq_b200")`.
"""
from . import _lib, design                     # noqa: F401
from .api import ComBat_seq                    # noqa: F401
from ._lib import CombatSeqError, device_count, kernel_launches   # noqa: F401
