"""B200-native bundle-adjustment solver: drop-in for the BA optimisation entry point of
MatteoWohlrapp/BundleAdjustment (ba_project/src/ba/Optimizer.{h,cpp}).  The product is the C-ABI shared library
(include/ba_b200.h, csrc/); this package is its ctypes binding plus the seeded synthetic scene generator.
Importing it without the built library raises -- there is no CPU fallback."""
from . import synthetic  # noqa: F401  (numpy only)
from . import capi       # noqa: F401  (raises ImportError when libba_b200.so is missing)
from . import sharding   # noqa: F401
from . import parity     # noqa: F401  (numpy only: the parity gate of bench.py / tools)
from .capi import Solver, default_options, solve_pose_only_batch, prune_frames  # noqa: F401

__all__ = ["synthetic", "capi", "sharding", "parity", "Solver", "default_options", "solve_pose_only_batch", "prune_frames"]
