"""B200-native ORB front-end: drop-in for ORB-SLAM2's ORBextractor / ORBmatcher hot path.

The product is liborbx.so (hand-written CUDA for sm_100a behind the C ABI in include/orbx.h).
This package is the Python host-side mirror of the reference's two classes, used by the parity
tests and bench.py. Nothing here computes on the CPU and nothing here imports oracle/.
"""
from ._capi import KP_DTYPE, LIB_PATH, OrbxError, SIGNATURES, lib, pinned_empty  # noqa: F401
from .extractor import (ORBextractor, blur7_u8, distribute, fast_atan2, resize_u8, sincosf,  # noqa: F401
                        sincosf_sweep)
from .matcher import Frame, ORBmatcher, Vocabulary, undistort_points  # noqa: F401
from . import sharding, synth  # noqa: F401


def device_count():
    return int(lib().orbx_device_count())
