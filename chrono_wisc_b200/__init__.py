"""chrono_wisc_b200 -- B200 (sm_100a) backend for the Chrono::Multicore granular-contact hot path.

Public surface: the reference-named system classes in ``system`` (host-side mirror of
ChSystemMulticoreSMC / ChSystemMulticoreNSC over the C ABI in include/chrono_b200.h) and the synthetic scenes
in ``scenes``.  The compute path is hand-written CUDA in ``csrc/``; there is no CPU fallback.
"""
from . import _lib, scenes, system  # noqa: F401
from .system import ChSystemMulticoreNSC, ChSystemMulticoreSMC, load_scene  # noqa: F401

__all__ = ["ChSystemMulticoreSMC", "ChSystemMulticoreNSC", "load_scene", "scenes", "system"]
