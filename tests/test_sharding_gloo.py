"""world_size-2 (and 3) CPU runs over gloo of the multi-GPU host logic: database row ranges, the all-gather
layout and the merge rule that orbx_match_merge_device implements on the GPU (restated here in numpy and
checked against the oracle's whole-database answer), plus frame sharding."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def merge_rule(idx, d1, d2):
    """Sequential top-2 over shards in ascending database order (R/src/ORBmatcher.cpp:282-291 applied to
    per-shard summaries): a strictly smaller minimum takes over and demotes the old best; an equal or larger
    one only competes for second place."""
    world, nq = idx.shape
    best = np.full(nq, np.iinfo(np.int32).max, np.int64)
    best2 = best.copy()
    bidx = np.full(nq, -1, np.int64)
    for s in range(world):
        a1, a2, ai = d1[s].astype(np.int64), d2[s].astype(np.int64), idx[s].astype(np.int64)
        take = a1 < best
        nb2 = np.where(take, np.minimum(best, a2), np.minimum(best2, a1))
        best2 = nb2
        bidx = np.where(take, ai, bidx)
        best = np.where(take, a1, best)
    return bidx.astype(np.int32), best.astype(np.int32), best2.astype(np.int32)


def _worker(rank, world, port, nq, ndb, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys
        from pathlib import Path
        root = Path(__file__).resolve().parents[1]
        sys.path[:0] = [str(root), str(root / "tests")]
        import oracle_lib
        import orbslam2_b200 as pkg
        db = pkg.synth.descriptors(ndb, 5)
        q = pkg.synth.planted_queries(db, nq, 6)
        db[ndb - 3] = db[7]          # duplicates across shards: the lower index must win, second == best
        q[1] = db[7]
        lo, hi = pkg.sharding.block_range(ndb, rank, world)
        idx, d1, d2, _ = oracle_lib.match_bf2(q, db[lo:hi], np.float32(0.6), 50, nthreads=2)
        part = torch.from_numpy(np.stack([idx + lo, d1, d2]).astype(np.int32))
        gathered = [torch.empty_like(part) for _ in range(world)]
        dist.all_gather(gathered, part)
        g = torch.stack(gathered).numpy()                       # [world, 3, nq]
        assert g.shape == (world, 3, nq)                        # what ncclAllGather delivers in csrc/orbx_shard.cu
        midx, m1, m2 = merge_rule(g[:, 0], g[:, 1], g[:, 2])
        eidx, e1, e2, _ = oracle_lib.match_bf2(q, db, np.float32(0.6), 50, nthreads=2)
        ok = np.array_equal(midx, eidx) and np.array_equal(m1, e1) and np.array_equal(m2, e2)
        ok = ok and midx[1] == 7 and m1[1] == 0 and m2[1] == 0
        # frame sharding: the ranks' blocks tile the sequence in order
        fr = torch.tensor(list(pkg.sharding.frame_range(1000, rank, world, group=2)))
        allfr = [torch.empty_like(fr) for _ in range(world)]
        dist.all_gather(allfr, fr)
        bounds = torch.stack(allfr).numpy()
        ok = ok and bounds[0, 0] == 0 and bounds[-1, 1] == 1000 and np.array_equal(bounds[1:, 0], bounds[:-1, 1])
        ok = ok and all(b % 2 == 0 for b in bounds.ravel())
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_match_merges_to_the_sequential_answer(world):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 64, 3000, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_block_range_properties(pkg):
    for n in (0, 1, 7, 10_000_000):
        for world in (1, 2, 3, 8):
            spans = [pkg.sharding.block_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        pkg.sharding.frame_range(7, 0, 2, group=2)
