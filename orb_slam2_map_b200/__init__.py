"""orb_slam2_map_b200 — B200-native (sm_100a) ORB front end behind the reference's operator API.

Product = liborbx.so (CUDA kernels + C ABI, include/orbx.h) and the C++ wrapper classes in
host/.  The Python modules here are ctypes plumbing for tests and bench.py.
"""
from . import capi, synth  # noqa: F401
from .capi import Extractor, OrbxError, KP_DTYPE  # noqa: F401
