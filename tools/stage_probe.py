"""Per-stage CUDA-event times of the device-resident pipeline vs frames per launch set."""
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
    for bs in (512, 128, 64, 32, 8, 1):
        ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=bs)
        ex.set_profiling(True)
        acc = {}
        reps = 0
        for it in range(6):
            for o in range(0, B, bs):
                ex.extract_batch_device(frames[o:].data_ptr(), bs, W, H, W, W * H, dk[o:].data_ptr(), dd[o:].data_ptr(), cap,
                                        dn[o:].data_ptr(), st)
                torch.cuda.synchronize()
                if it >= 2:
                    for k, v in ex.last_stage_ms().items():
                        acc[k] = acc.get(k, 0.0) + v
                    reps += 1
                if o >= 8 * bs:
                    break
        print(f"batch {bs:4d}: " + "  ".join(f"{k} {v / reps * 1e3:7.1f}us" for k, v in acc.items()), flush=True)
        ex.close()
