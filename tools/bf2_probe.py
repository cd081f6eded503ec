#!/usr/bin/env python3
"""Brute-force top-2 timing on one GPU for config 5 and for one of its 8 database shards (10k queries x 10M / 1.25M
rows, device-resident, CUDA events). Usage: python tools/bf2_probe.py   (ORBX_LIB selects an experiment build)"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import orbslam2_b200 as pkg

torch.cuda.set_device(0)
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    m = pkg.ORBmatcher(0.6)
    for nq, ndb in ((10_000, 10_000_000), (10_000, 1_250_000), (2048, 1_000_000), (2000, 20_000), (10_000, 3_000_000)):
        q = torch.randint(0, 256, (nq, 32), generator=g, device="cuda", dtype=torch.uint8)
        out = torch.empty((3, nq), dtype=torch.int32, device="cuda")
        db = torch.randint(0, 256, (ndb, 32), generator=g, device="cuda", dtype=torch.uint8)
        s = st.cuda_stream
        for _ in range(2):
            m.match_bf2_device(q.data_ptr(), nq, db.data_ptr(), ndb, 0, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), s)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            m.match_bf2_device(q.data_ptr(), nq, db.data_ptr(), ndb, 0, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"nq {nq} ndb {ndb}: {ms:.3f} ms  {nq * ndb / ms / 1e6:.1f} Gpairs/s  checksum {int(out.sum().item())}")
        del db
