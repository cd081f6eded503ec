"""Device-resident extraction rate vs batch size (one stream, back-to-back launches): how much of the host-buffer
path's gap to the 512-frame device rate is the smaller launches of 32-frame chunks?"""
import importlib
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
pkg = importlib.import_module("orb-slam2-_b200")
W, H, B = 640, 480, 512
base = np.stack([pkg.synth.frame(W, H, 1000 + i) for i in range(16)])
frames = torch.from_numpy(base).cuda().repeat(B // 16, 1, 1).contiguous()
cap = 1064
dk = torch.empty((B, cap, 7), dtype=torch.float32, device="cuda")
dd = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
dn = torch.zeros(B, dtype=torch.int32, device="cuda")
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    st = torch.cuda.current_stream().cuda_stream
    for bs in (512, 256, 128, 64, 32, 16):
        ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=bs)
        ex.set_profiling(False)

        def run():
            for o in range(0, B, bs):
                ex.extract_batch_device(frames[o:].data_ptr(), bs, W, H, W, W * H, dk[o:].data_ptr(), dd[o:].data_ptr(), cap,
                                        dn[o:].data_ptr(), st)
        for _ in range(3):
            run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"batch {bs:4d}: {ms:.3f} ms per {B} frames  ({B / ms * 1e3:.0f} frames/s)", flush=True)
