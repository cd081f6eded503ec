"""ctypes binding of liborbx.so (include/orbx.h). Fails loudly when the CUDA library is missing:
there is no CPU fallback anywhere in this package."""
import ctypes as C
import os
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
# ORBX_LIB: an experiment build of the same library (orb-slam2-_b200/build.py build_variant); never a different backend
LIB_PATH = Path(os.environ["ORBX_LIB"]) if os.environ.get("ORBX_LIB") else HERE / "liborbx.so"

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"orbx error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("device", C.c_int32),
                ("max_width", C.c_int32), ("max_height", C.c_int32), ("max_batch", C.c_int32)]


_vp, _i32, _i64, _f32, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t

# name -> (restype, argtypes); every symbol include/orbx.h declares
SIGNATURES = {
    "orbx_last_error": (C.c_char_p, []),
    "orbx_version": (C.c_char_p, []),
    "orbx_device_count": (C.c_int, []),
    "orbx_host_alloc": (C.c_int, [_i32, _sz, C.POINTER(_vp)]),
    "orbx_host_free": (C.c_int, [_vp]),
    "orbx_extractor_create": (C.c_int, [C.POINTER(Params), C.POINTER(_vp)]),
    "orbx_extractor_destroy": (None, [_vp]),
    "orbx_extractor_tables": (C.c_int, [_vp] + [_vp] * 6),
    "orbx_extractor_set_pipeline": (C.c_int, [_vp, _i32, _i32]),
    "orbx_extract": (C.c_int, [_vp, _vp, _i32, _i32, _sz, _vp, _vp, _i32, C.POINTER(_i32)]),
    "orbx_extract_batch": (C.c_int, [_vp, C.POINTER(_vp), _i32, _i32, _i32, _sz, _vp, _vp, _i32, _vp]),
    "orbx_extract_batch_submit": (C.c_int, [_vp, C.POINTER(_vp), _i32, _i32, _i32, _sz, _vp, _vp, _i32, _vp, C.POINTER(_i32)]),
    "orbx_stereo_batch_submit": (C.c_int, [_vp, C.POINTER(_vp), _i32, _i32, _i32, _sz, _vp, _vp, _i32, _vp, _f32, _f32, _vp, _vp, _vp,
                                           C.POINTER(_i32)]),
    "orbx_extract_batch_wait": (C.c_int, [_vp, _i32]),
    "orbx_extract_batch_device": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _sz, _sz, _vp, _vp, _i32, _vp, _vp]),
    "orbx_level_size": (C.c_int, [_vp, _i32, C.POINTER(_i32), C.POINTER(_i32)]),
    "orbx_get_level": (C.c_int, [_vp, _i32, _i32, _vp, _sz]),
    "orbx_get_pyramid": (C.c_int, [_vp, _i32, _vp, _sz, _vp, _vp, _vp]),
    "orbx_get_blurred_level": (C.c_int, [_vp, _i32, _i32, _vp, _sz]),
    "orbx_get_level_candidates": (C.c_int, [_vp, _i32, _i32, _vp, _i32, C.POINTER(_i32)]),
    "orbx_launch_count": (_i64, [_vp]),
    "orbx_set_profiling": (C.c_int, [_vp, _i32]),
    "orbx_last_stage_ms": (C.c_int, [_vp, _vp]),
    "orbx_resize_u8": (C.c_int, [_i32, _vp, _i32, _i32, _sz, _vp, _i32, _i32, _sz]),
    "orbx_blur7_u8": (C.c_int, [_i32, _vp, _i32, _i32, _sz, _vp, _sz]),
    "orbx_fast_atan2": (C.c_int, [_i32, _vp, _vp, _vp, _i64]),
    "orbx_sincosf": (C.c_int, [_i32, _vp, _vp, _vp, _i64]),
    "orbx_sincosf_sweep": (C.c_int, [_i32, C.c_uint32, C.c_uint32, _vp, _vp]),
    "orbx_distribute": (C.c_int, [_i32, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _i32,
                                  C.POINTER(_i32)]),
    "orbx_matcher_create": (C.c_int, [_i32, C.POINTER(_vp)]),
    "orbx_matcher_destroy": (None, [_vp]),
    "orbx_matcher_launch_count": (_i64, [_vp]),
    "orbx_descriptor_distance": (C.c_int, [_vp, _vp, _vp, _vp, _i64]),
    "orbx_match_bf2": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _f32, _i32, _vp, _vp, _vp, _vp]),
    "orbx_match_bf2_device": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "orbx_match_bf2_noalloc": (C.c_int, [_i32, _vp, _i32, _vp, _i64, _f32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "orbx_match_window": (C.c_int, [_vp, _vp, _vp, _i32, _f32, _f32, _f32, _f32, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp,
                                    _i32, C.POINTER(_i32)]),
    "orbx_stereo_match": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _vp, _vp, _i32, _f32, _f32, _vp, _vp,
                                    C.POINTER(_i32)]),
    "orbx_stereo_match_batch_device": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _i32, _f32, _f32, _vp, _vp, _vp,
                                                 _vp]),
    "orbx_match_merge_device": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _f32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "orbx_match_blocks": (C.c_int, [_vp, _vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i64, C.POINTER(_i64)]),
    "orbx_distinctive": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _vp]),
    "orbx_undistort_points": (C.c_int, [_i32, _vp, _i64, _vp, _vp, _i32, _vp]),
    "orbx_frame_create": (C.c_int, [_vp, C.POINTER(_vp)]),
    "orbx_frame_destroy": (None, [_vp]),
    "orbx_frame_set": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _vp]),
    "orbx_frame_get_keys_un": (C.c_int, [_vp, _vp]),
    "orbx_frame_match_window": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i32, C.POINTER(_i32)]),
    "orbx_vocabulary_create": (C.c_int, [_vp, _i32, _vp, _vp, _i32, C.POINTER(_vp)]),
    "orbx_vocabulary_destroy": (None, [_vp]),
    "orbx_vocabulary_descend": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _vp]),
    "orbx_shard_unique_id": (C.c_int, [_vp]),
    "orbx_shard_group_create": (C.c_int, [_vp, _i32, C.POINTER(_vp)]),
    "orbx_shard_group_create_rank": (C.c_int, [_i32, _i32, _i32, _vp, C.POINTER(_vp)]),
    "orbx_shard_group_destroy": (None, [_vp]),
    "orbx_shard_group_info": (C.c_int, [_vp, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32)]),
    "orbx_shard_launch_count": (_i64, [_vp]),
    "orbx_shard_block_range": (C.c_int, [_i64, _i32, _i32, _i32, C.POINTER(_i64), C.POINTER(_i64)]),
    "orbx_shard_set_database": (C.c_int, [_vp, _vp, _i64]),
    "orbx_shard_set_database_device": (C.c_int, [_vp, _i32, _vp, _i64, _i64, _i64]),
    "orbx_shard_match_bf2": (C.c_int, [_vp, _vp, _i32, _f32, _i32, _vp, _vp, _vp, _vp]),
    "orbx_shard_match_bf2_device": (C.c_int, [_vp, _i32, _vp, _i32, _f32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "orbx_shard_extract_sequence": (C.c_int, [_vp, C.POINTER(Params), C.POINTER(_vp), _i64, _i32, _i32, _i32, _sz, _vp, _vp, _i32,
                                              _vp, _f32, _f32, _vp, _vp, _vp]),
    "orbx_distinctive_device": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _vp, _vp]),
}

_lib = None


def lib():
    """The loaded liborbx.so. Raises if it has not been built (python orb-slam2-_b200/build.py)."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise OrbxError(-2, f"{LIB_PATH} is missing: build it with __graft_entry__.build(); "
                                "there is no CPU fallback")
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise OrbxError(rc, lib().orbx_last_error().decode(errors="replace"))


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _host_free(address):
    lib().orbx_host_free(C.c_void_p(address))


def pinned_empty(shape, dtype, device=0):
    """numpy array in page-locked host memory (orbx_host_alloc): frames and results in such arrays make the
    copies of orbx_extract_batch asynchronous and let results land in them without a staging copy.
    The block is freed when the last view of the array is gone."""
    import weakref

    import numpy as np
    dt = np.dtype(dtype)
    shape = (shape,) if isinstance(shape, int) else tuple(shape)
    n = int(np.prod(shape)) * dt.itemsize
    p = C.c_void_p()
    check(lib().orbx_host_alloc(device, max(n, 1), C.byref(p)))
    owner = np.frombuffer((C.c_uint8 * max(n, 1)).from_address(p.value), dtype=np.uint8)  # every view's .base
    weakref.finalize(owner, _host_free, p.value)
    return owner[:n].view(dt).reshape(shape)
