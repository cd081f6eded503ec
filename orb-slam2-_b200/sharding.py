"""How the hot path shards across the GPUs of one box. One process per GPU.

* extraction: frames are independent -> contiguous blocks of frame indices per rank (sequence order is kept
  inside a rank, a stereo pair stays on one rank), no data-path collective.
* brute-force Hamming: the descriptor database is cut into `world` contiguous row ranges (global index order
  is preserved, so "lowest index wins" survives the merge); queries are replicated; each rank produces
  (idx, d1, d2) per query; ONE all-gather; every rank merges with orbx_match_merge_device.
"""


def block_range(n, rank, world):
    """[lo, hi) of `n` items owned by `rank`: contiguous, sizes differ by at most one, covers 0..n."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return n * rank // world, n * (rank + 1) // world


def frame_range(nframes, rank, world, group=1):
    """Frames of a sequence owned by `rank`; `group` frames stay together (2 for stereo pairs)."""
    if nframes % group:
        raise ValueError("sequence length must be a multiple of the group size")
    lo, hi = block_range(nframes // group, rank, world)
    return lo * group, hi * group


def gather_layout(world, nq):
    """Shape of the all-gathered partial results handed to orbx_match_merge_device: for each of idx, d1, d2
    a [world][nq] int32 array, shards in ascending database order."""
    return (world, nq)
