"""How the hot path shards across the GPUs of one box: the Python mirror of the orbx_shard_* C ABI
(csrc/orbx_shard.cu). All the logic -- block ranges, NCCL communicator, all-gather, merge -- lives in liborbx.so; this
file only marshals arguments.

* extraction: frames are independent -> contiguous blocks of frame indices per rank (sequence order is kept
  inside a rank, a stereo pair stays on one rank), no data-path collective.
* brute-force Hamming: the descriptor database is cut into contiguous row ranges (global index order is preserved,
  so "lowest index wins" survives the merge); queries are replicated; each rank produces (idx, d1, d2) per query;
  ONE ncclAllGather; every rank merges.
"""
import ctypes as C

import numpy as np

from ._capi import KP_DTYPE, Params, check, lib, ptr


def block_range(n, rank, world, group=1):
    """[lo, hi) of `n` items owned by `rank` (orbx_shard_block_range): contiguous blocks of whole groups."""
    lo, hi = C.c_int64(), C.c_int64()
    rc = lib().orbx_shard_block_range(n, rank, world, group, C.byref(lo), C.byref(hi))
    if rc:
        raise ValueError("bad rank/world")
    return lo.value, hi.value


def frame_range(nframes, rank, world, group=1):
    """Frames of a sequence owned by `rank`; `group` frames stay together (2 for stereo pairs)."""
    if nframes % group:
        raise ValueError("sequence length must be a multiple of the group size")
    return block_range(nframes, rank, world, group)


def unique_id():
    """128-byte NCCL id for ShardGroup.from_rank (rank 0 makes it, the launcher broadcasts it)."""
    buf = np.zeros(128, np.uint8)
    check(lib().orbx_shard_unique_id(ptr(buf)))
    return buf


class ShardGroup:
    """orbx_shard_group: every rank of one process (devices=[...]) or one rank of a multi-process job (from_rank)."""

    def __init__(self, devices=None, _handle=None):
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
            return
        dev = np.ascontiguousarray(devices, np.int32)
        check(lib().orbx_shard_group_create(ptr(dev), len(dev), C.byref(self._h)))

    @classmethod
    def from_rank(cls, device, rank, nranks, uid):
        h = C.c_void_p()
        uid = np.ascontiguousarray(uid, np.uint8) if uid is not None else None
        check(lib().orbx_shard_group_create_rank(device, rank, nranks, ptr(uid) if uid is not None else None, C.byref(h)))
        return cls(_handle=h)

    def close(self):
        if self._h:
            lib().orbx_shard_group_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        a, b, c, d = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        check(lib().orbx_shard_group_info(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return {"nranks": a.value, "nlocal": b.value, "first_rank": c.value, "nccl_version": d.value}

    def launch_count(self):
        return int(lib().orbx_shard_launch_count(self._h))

    def set_database(self, db):
        db = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
        check(lib().orbx_shard_set_database(self._h, ptr(db), db.shape[0]))

    def set_database_device(self, local_index, d_rows_ptr, row_lo, row_hi, ndb):
        check(lib().orbx_shard_set_database_device(self._h, local_index, d_rows_ptr, row_lo, row_hi, ndb))

    def match_bf2(self, q, ratio=0.6, th=50):
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
        nq = q.shape[0]
        idx, d1, d2 = (np.empty(nq, np.int32) for _ in range(3))
        acc = np.empty(nq, np.uint8)
        check(lib().orbx_shard_match_bf2(self._h, ptr(q), nq, ratio, th, ptr(idx), ptr(d1), ptr(d2), ptr(acc)))
        return idx, d1, d2, acc

    def match_bf2_device(self, local_index, d_q, nq, ratio, th, d_idx, d_d1, d_d2, d_acc, stream=0):
        check(lib().orbx_shard_match_bf2_device(self._h, local_index, d_q, nq, ratio, th, d_idx, d_d1, d_d2, d_acc, stream))

    def extract_sequence(self, params, frames, kps, desc, n, group=1, stereo=None, total=None, block=None):
        """orbx_shard_extract_sequence on a (count, H, W) uint8 array. params: dict of orbx_params fields (device is
        replaced per rank). stereo = (mbf, mb, u_right, depth, n_matched) for L R L R ... sequences (group 2).
        Multi-process jobs: total = length of the whole sequence and block = (lo, hi) = this process's block
        (orbx_shard_block_range); the arrays then hold only that block (the C entry point indexes its arguments by
        global frame number, so it is handed base addresses moved back by lo rows; rows outside the block are never
        touched)."""
        if frames.dtype != np.uint8 or frames.ndim != 3 or frames.strides[2] != 1:
            raise TypeError("frames must be a (count, H, W) uint8 array with contiguous rows")
        cnt, h, w = frames.shape
        cap = kps.shape[1]
        if kps.dtype != KP_DTYPE or kps.shape[0] < cnt or desc.shape[:2] != kps.shape[:2] or n.shape[0] < cnt or n.dtype != np.int32:
            raise TypeError("output arrays do not match the sequence")
        lo = 0
        if total is None:
            total = cnt
        else:
            lo, hi = block
            if hi - lo != cnt or lo % group:
                raise ValueError("block does not match the arrays")
        p = Params(params["nfeatures"], params["scale_factor"], params["nlevels"], params["ini_th_fast"],
                   params["min_th_fast"], 0, params.get("max_width", w), params.get("max_height", h),
                   params.get("max_batch", 256))
        ptrs = np.zeros(total, np.int64)
        ptrs[lo:lo + cnt] = frames.ctypes.data + frames.strides[0] * np.arange(cnt, dtype=np.int64)
        mbf, mb, ur, dp, nm = stereo if stereo else (0.0, 0.0, None, None, None)

        def base(a, row_bytes, first):
            return C.c_void_p(a.ctypes.data - first * row_bytes) if a is not None else None
        check(lib().orbx_shard_extract_sequence(self._h, C.byref(p), C.cast(ptr(ptrs), C.POINTER(C.c_void_p)), total, group, w, h,
                                                frames.strides[1], base(kps, cap * KP_DTYPE.itemsize, lo),
                                                base(desc, cap * 32, lo), cap, base(n, 4, lo), mbf, mb,
                                                base(ur, cap * 4, lo // 2), base(dp, cap * 4, lo // 2), base(nm, 4, lo // 2)))
