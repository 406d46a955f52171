"""qflex_b200 -- B200-native tensor-network contraction engine with qFlex's surface.

Python surface (mirrors the reference's pybind module ``qflex``,
src/pybind_main.cpp:75-138):

    import qflex_b200 as qflex
    qflex.simulate({'circuit_filename': ..., 'ordering_filename': ...,
                    'grid_filename': ..., 'final_state': ...})

The numeric path is hand-written sm_100a CUDA behind the C ABI in
include/qflex_b200.h; there is no CPU fallback.
"""
from ._host import (HostLibrary, Prepared, QflexError, default_library,  # noqa: F401
                    simulate, restrict_cuts, trim_device_cache, STAT_NAMES)
from . import _device as device  # noqa: F401

__all__ = ["simulate", "HostLibrary", "Prepared", "QflexError", "default_library",
           "device", "STAT_NAMES", "restrict_cuts", "trim_device_cache"]
