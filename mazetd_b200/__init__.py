"""mazetd_b200 -- B200-native (sm_100a) replacement for mazetd's flow-field hot path.

The product is the C-ABI shared library ``libmazetd_b200.so`` (include/mazetd_b200.h, mazetd_b200/csrc/).
``capi`` binds it with ctypes; ``host`` mirrors the reference's map interface (Grid / Directions / Waypoints,
sources/grid.h) on top of it; ``synth`` produces the fixed-seed synthetic inputs of the bench configs.
Nothing in this package touches oracle/: the CUDA path is the only path.
"""
from . import synth  # noqa: F401

__all__ = ["capi", "host", "synth", "build"]
