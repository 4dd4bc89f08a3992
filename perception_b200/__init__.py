"""perception_b200 -- B200-native (sm_100a) implementation of PERCH 2.0's render + cost hot path.

The compute lives in ``libperch_b200.so`` (CUDA, C-ABI in ``include/perch_b200.h``).  This package is the thin
Python host mirror of the reference's ``cuda_renderer`` interface for the repo's Python experiment scripts
(reference: sbpl_perception/src/scripts/tools/fat_dataset/perch.py drives the C++ binary; here the library is
loaded directly).  There is no CPU fallback: importing :mod:`perception_b200.renderer` without the built library,
or creating a context without a B200, raises.
"""
from ._capi import LIB_PATH, PerchError, load_library  # noqa: F401

__all__ = ["LIB_PATH", "PerchError", "load_library"]
