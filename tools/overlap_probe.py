"""Do the pipeline's kernels overlap usefully across streams? 2048 device-resident frames processed (a) by one handle on
one stream in 512-frame launch sets, (b) by S handles on S streams, each taking every S-th launch set of 512/S... frames.
FAST is ALU-bound, blur / describe / octree are not: kernels of different sets may fill each other's idle pipes."""
import importlib
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
pkg = importlib.import_module("orb-slam2-_b200")
W, H, B = 640, 480, 2048
base = np.stack([pkg.synth.frame(W, H, 1000 + i) for i in range(32)])
frames = torch.from_numpy(base).cuda().repeat(B // 32, 1, 1).contiguous()
cap = 1064
dk = torch.empty((B, cap, 7), dtype=torch.float32, device="cuda")
dd = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
dn = torch.zeros(B, dtype=torch.int32, device="cuda")

for nstreams, bs in ((1, 512), (2, 512), (2, 256), (3, 256), (4, 256), (4, 128), (2, 128), (8, 64)):
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    exs = [pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=bs) for _ in range(nstreams)]

    def run():
        for k, o in enumerate(range(0, B, bs)):
            s = k % nstreams
            exs[s].extract_batch_device(frames[o:].data_ptr(), bs, W, H, W, W * H, dk[o:].data_ptr(), dd[o:].data_ptr(), cap,
                                        dn[o:].data_ptr(), streams[s].cuda_stream)
    for _ in range(2):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams:
        s.wait_event(e0)
    for _ in range(4):
        run()
    for s in streams:
        e = torch.cuda.Event()
        e.record(s)
        torch.cuda.current_stream().wait_event(e)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 4
    print(f"{nstreams} stream(s) x {bs:3d}-frame sets: {ms:.3f} ms per {B} frames  ({B / ms * 1e3:.0f} frames/s)", flush=True)
    for ex in exs:
        ex.close()
