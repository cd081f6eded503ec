"""Python mirror of ORB_SLAM2::ORBextractor (R/include/ORBextractor.h:45-123) over the C ABI.

Same constructor arguments, same getters, `__call__(image)` for operator() and a lazily filled
`mvImagePyramid`. All compute happens in liborbx.so (CUDA, sm_100a); numpy only carries buffers.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import KP_DTYPE, Params, check, lib, ptr


class ORBextractor:
    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, *, device=0,
                 max_width=1920, max_height=1080, max_batch=1):
        self.nfeatures, self.nlevels = int(nfeatures), int(nlevels)
        self.scaleFactor = float(np.float32(scaleFactor))
        self.iniThFAST, self.minThFAST = int(iniThFAST), int(minThFAST)
        self.device, self.max_batch = device, max_batch
        p = Params(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, max_width, max_height, max_batch)
        h = C.c_void_p()
        check(lib().orbx_extractor_create(C.byref(p), C.byref(h)))
        self._h = h
        n = nlevels
        self._scale, self._inv_scale, self._sigma2, self._inv_sigma2 = (np.empty(n, np.float32) for _ in range(4))
        self.mnFeaturesPerLevel = np.empty(n, np.int32)
        self.umax = np.empty(16, np.int32)
        check(lib().orbx_extractor_tables(self._h, ptr(self._scale), ptr(self._inv_scale), ptr(self._sigma2),
                                          ptr(self._inv_sigma2), ptr(self.mnFeaturesPerLevel), ptr(self.umax)))
        self._out_cap = 0

    def close(self):
        if getattr(self, "_h", None):
            lib().orbx_extractor_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- getters (R/include/ORBextractor.h:66-86) ----
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self.scaleFactor

    def GetScaleFactors(self):
        return self._scale.copy()

    def GetInverseScaleFactors(self):
        return self._inv_scale.copy()

    def GetScaleSigmaSquares(self):
        return self._sigma2.copy()

    def GetInverseScaleSigmaSquares(self):
        return self._inv_sigma2.copy()

    def _cap(self):
        return int(self.nfeatures + 8 * self.nlevels + 512)

    # ---- operator() ----
    def __call__(self, image, mask=None):
        """(keypoints, descriptors) for one CV_8UC1 image. Empty image -> (empty, None), like the
        reference's early return / descriptors.release() (R/src/ORBextractor.cpp:1123, :1145)."""
        image = np.asarray(image)
        if image.size == 0:
            return np.empty(0, KP_DTYPE), None
        if image.dtype != np.uint8 or image.ndim != 2 or image.strides[1] != 1:
            raise TypeError("image must be a CV_8UC1 (uint8, 2-D, row-contiguous) array")
        cap = self._cap()
        kps = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n = C.c_int32(0)
        rc = lib().orbx_extract(self._h, ptr(image), image.shape[1], image.shape[0], image.strides[0],
                                ptr(kps), ptr(desc), cap, C.byref(n))
        check(rc)
        if n.value == 0:
            return np.empty(0, KP_DTYPE), None
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images):
        """List of same-shape images -> list of (keypoints, descriptors)."""
        imgs = [np.asarray(im) for im in images]
        h, w = imgs[0].shape
        stride = imgs[0].strides[0]
        for im in imgs:
            if im.dtype != np.uint8 or im.shape != (h, w) or im.strides != (stride, 1):
                raise TypeError("batch images must share shape, stride and dtype uint8")
        cnt, cap = len(imgs), self._cap()
        arr = (C.c_void_p * cnt)(*[im.ctypes.data for im in imgs])
        kps = np.empty((cnt, cap), KP_DTYPE)
        desc = np.empty((cnt, cap, 32), np.uint8)
        n = np.zeros(cnt, np.int32)
        check(lib().orbx_extract_batch(self._h, arr, cnt, w, h, stride, ptr(kps), ptr(desc), cap, ptr(n)))
        return [(kps[f, :n[f]].copy(), desc[f, :n[f]].copy()) for f in range(cnt)]

    def set_pipeline(self, chunk_frames=0, lanes=0):
        """Frames per chunk / chunks in flight of the host-buffer batch path (0 = library default)."""
        check(lib().orbx_extractor_set_pipeline(self._h, chunk_frames, lanes))

    def extract_batch_into(self, frames, kps, desc, n):
        """orbx_extract_batch on a (B, H, W) uint8 host array (ideally pinned) into caller-owned outputs:
        kps (B, cap) KP_DTYPE, desc (B, cap, 32) uint8, n (B,) int32. No per-frame Python work: this is the
        call bench.py times end to end."""
        if frames.dtype != np.uint8 or frames.ndim != 3 or frames.strides[2] != 1:
            raise TypeError("frames must be a (B, H, W) uint8 array with contiguous rows")
        cnt, h, w = frames.shape
        cap = kps.shape[1]
        if kps.dtype != KP_DTYPE or kps.shape[0] < cnt or desc.shape[:2] != kps.shape[:2] or n.shape[0] < cnt:
            raise TypeError("output arrays do not match the batch")
        ptrs = frames.ctypes.data + frames.strides[0] * np.arange(cnt, dtype=np.int64)   # imgs[f], no Python loop
        check(lib().orbx_extract_batch(self._h, C.cast(ptr(ptrs), C.POINTER(C.c_void_p)), cnt, w, h, frames.strides[1],
                                       ptr(kps), ptr(desc), cap, ptr(n)))

    def extract_batch_submit(self, frames, kps, desc, n):
        """orbx_extract_batch_submit: the streaming form of extract_batch_into. All four arrays must be page-locked
        (pinned_empty) and stay untouched until extract_batch_wait(ticket). Returns the ticket."""
        if frames.dtype != np.uint8 or frames.ndim != 3 or frames.strides[2] != 1:
            raise TypeError("frames must be a (B, H, W) uint8 array with contiguous rows")
        cnt, h, w = frames.shape
        cap = kps.shape[1]
        if kps.dtype != KP_DTYPE or kps.shape[0] < cnt or desc.shape[:2] != kps.shape[:2] or n.shape[0] < cnt or n.dtype != np.int32:
            raise TypeError("output arrays do not match the batch")
        ptrs = frames.ctypes.data + frames.strides[0] * np.arange(cnt, dtype=np.int64)
        ticket = C.c_int32(0)
        check(lib().orbx_extract_batch_submit(self._h, C.cast(ptr(ptrs), C.POINTER(C.c_void_p)), cnt, w, h, frames.strides[1],
                                              ptr(kps), ptr(desc), cap, ptr(n), C.byref(ticket)))
        return ticket.value

    def stereo_batch_submit(self, frames, kps, desc, n, mbf, mb, u_right, depth, n_matched):
        """orbx_stereo_batch_submit: frames are L R L R ...; every chunk also runs Frame::ComputeStereoMatches on its
        pairs. u_right / depth (B/2, cap) float32 and n_matched (B/2,) int32, all page-locked. Returns the ticket."""
        if frames.dtype != np.uint8 or frames.ndim != 3 or frames.strides[2] != 1:
            raise TypeError("frames must be a (B, H, W) uint8 array with contiguous rows")
        cnt, h, w = frames.shape
        cap = kps.shape[1]
        if (kps.dtype != KP_DTYPE or kps.shape[0] < cnt or desc.shape[:2] != kps.shape[:2] or n.shape[0] < cnt or n.dtype != np.int32
                or u_right.dtype != np.float32 or depth.dtype != np.float32 or n_matched.dtype != np.int32
                or u_right.shape != (cnt // 2, cap) or depth.shape != (cnt // 2, cap) or n_matched.shape[0] < cnt // 2):
            raise TypeError("output arrays do not match the batch")
        ptrs = frames.ctypes.data + frames.strides[0] * np.arange(cnt, dtype=np.int64)
        ticket = C.c_int32(0)
        check(lib().orbx_stereo_batch_submit(self._h, C.cast(ptr(ptrs), C.POINTER(C.c_void_p)), cnt, w, h, frames.strides[1],
                                             ptr(kps), ptr(desc), cap, ptr(n), mbf, mb, ptr(u_right), ptr(depth),
                                             ptr(n_matched), C.byref(ticket)))
        return ticket.value

    def extract_batch_wait(self, ticket):
        check(lib().orbx_extract_batch_wait(self._h, ticket))

    def extract_batch_device(self, d_imgs_ptr, count, width, height, row_stride, frame_stride, d_kps_ptr,
                             d_desc_ptr, cap, d_n_ptr, stream=0):
        """Device-resident batch (raw device pointers, e.g. torch tensors' data_ptr()); asynchronous."""
        check(lib().orbx_extract_batch_device(self._h, d_imgs_ptr, count, width, height, row_stride, frame_stride,
                                              d_kps_ptr, d_desc_ptr, cap, d_n_ptr, stream))

    # ---- mvImagePyramid (R/include/ORBextractor.h:88), copied back on demand ----
    def level_size(self, level):
        w, h = C.c_int32(), C.c_int32()
        check(lib().orbx_level_size(self._h, level, C.byref(w), C.byref(h)))
        return w.value, h.value

    def image_pyramid_level(self, level, frame=0, blurred=False):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        fn = lib().orbx_get_blurred_level if blurred else lib().orbx_get_level
        check(fn(self._h, frame, level, ptr(out), w))
        return out

    def image_pyramid(self, frame=0):
        """Levels 1 .. nlevels-1 of `frame` of the last batch in one batch of copies (orbx_get_pyramid): a list of
        (h, w) views into one buffer, row stride = the device pitch. Entry 0 is None (level 0 is the caller's image)."""
        off = (C.c_size_t * 16)()
        pitch = (C.c_int32 * 16)()
        nbytes = C.c_size_t()
        check(lib().orbx_get_pyramid(self._h, frame, None, 0, off, pitch, C.byref(nbytes)))
        buf = np.empty(max(nbytes.value, 1), np.uint8)
        check(lib().orbx_get_pyramid(self._h, frame, ptr(buf), buf.size, off, pitch, C.byref(nbytes)))
        out = [None]
        for l in range(1, self.nlevels):
            w, h = self.level_size(l)
            out.append(np.lib.stride_tricks.as_strided(buf[off[l]:], shape=(h, w), strides=(pitch[l], 1), writeable=False))
        return out

    @property
    def mvImagePyramid(self):
        return [self.image_pyramid_level(l) for l in range(self.nlevels)]

    # ---- introspection for tests / bench ----
    def level_candidates(self, level, frame=0):
        n = C.c_int32()
        check(lib().orbx_get_level_candidates(self._h, frame, level, None, 0, C.byref(n)))
        out = np.empty((max(n.value, 1), 3), np.int32)
        check(lib().orbx_get_level_candidates(self._h, frame, level, ptr(out), n.value, C.byref(n)))
        return out[:n.value]

    def launch_count(self):
        return int(lib().orbx_launch_count(self._h))

    def set_profiling(self, on):
        check(lib().orbx_set_profiling(self._h, int(on)))

    def last_stage_ms(self):
        ms = np.zeros(6, np.float32)
        check(lib().orbx_last_stage_ms(self._h, ptr(ms)))
        return dict(zip(["pyramid", "fast", "octree", "blur", "describe", "total"], ms.tolist()))


# ---- stand-alone primitives (kernel-level parity tests) ------------------------------------------
def resize_u8(src, dw, dh, device=0):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    check(lib().orbx_resize_u8(device, ptr(src), src.shape[1], src.shape[0], src.strides[0], ptr(dst), dw, dh, dw))
    return dst


def blur7_u8(src, device=0):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty_like(src)
    check(lib().orbx_blur7_u8(device, ptr(src), src.shape[1], src.shape[0], src.strides[0], ptr(dst), src.shape[1]))
    return dst


def fast_atan2(y, x, device=0):
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(y)
    check(lib().orbx_fast_atan2(device, ptr(y), ptr(x), ptr(out), y.size))
    return out


def sincosf(a, device=0):
    a = np.ascontiguousarray(a, np.float32)
    s, c = np.empty_like(a), np.empty_like(a)
    check(lib().orbx_sincosf(device, ptr(a), ptr(s), ptr(c), a.size))
    return s, c


def sincosf_sweep(lo_bits, hi_bits, device=0):
    n = hi_bits - lo_bits
    s, c = np.empty(n, np.float32), np.empty(n, np.float32)
    check(lib().orbx_sincosf_sweep(device, lo_bits, hi_bits, ptr(s), ptr(c)))
    return s, c


def distribute(xys, min_x, max_x, min_y, max_y, n_features, cell_w, cell_h, device=0):
    xys = np.ascontiguousarray(xys, np.int32).reshape(-1, 3)
    cap = n_features + 512
    out = np.empty((cap, 3), np.int32)
    n = C.c_int32()
    check(lib().orbx_distribute(device, ptr(xys), len(xys), min_x, max_x, min_y, max_y, n_features, cell_w, cell_h,
                                ptr(out), cap, C.byref(n)))
    return out[:n.value].copy()
