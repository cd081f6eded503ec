"""Where does the host-buffer path (orbx_extract_batch) spend its time? Sweeps chunk size x lanes for the e2e call and
measures the device-resident rate at the same batch sizes (small batches fill the GPU less well)."""
import importlib
import os
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def one(lanes, chunk, B=512):
    os.environ["ORBX_LANES"] = str(lanes)
    if chunk > 0:
        os.environ["ORBX_CHUNK"] = str(chunk)   # 0 = the library default
    pkg = importlib.import_module("orb-slam2-_b200")
    W, H = 640, 480
    frames = pkg.pinned_empty((B, H, W), np.uint8)
    base = [pkg.synth.frame(W, H, 1000 + i) for i in range(16)]
    for i in range(B):
        frames[i] = np.roll(base[i % 16], i // 16, axis=1)
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    cap = 1064
    k, d, n = pkg.pinned_empty((B, cap), pkg.KP_DTYPE), pkg.pinned_empty((B, cap, 32), np.uint8), np.zeros(B, np.int32)
    for _ in range(3):
        ex.extract_batch_into(frames, k, d, n)
    t0 = time.perf_counter()
    reps = 8
    for _ in range(reps):
        ex.extract_batch_into(frames, k, d, n)
    dt = (time.perf_counter() - t0) / reps
    print(f"lanes {lanes} chunk {chunk:4d}: {B / dt:9.0f} frames/s  ({dt * 1e3:.2f} ms per {B} frames)", flush=True)


if __name__ == "__main__":
    if len(sys.argv) == 3:
        one(int(sys.argv[1]), int(sys.argv[2]))
    else:
        for lanes, chunk in [(4, 0), (1, 512), (2, 256), (4, 128), (3, 64), (4, 64), (6, 64), (4, 32), (8, 32), (8, 16)]:
            subprocess.run([sys.executable, __file__, str(lanes), str(chunk)], check=False)
