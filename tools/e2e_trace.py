"""One traced orbx_extract_batch call (ORBX_TRACE=1): per-chunk device timeline on stderr."""
import importlib
import os
import sys
from pathlib import Path

import numpy as np

os.environ["ORBX_TRACE"] = "1"
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
pkg = importlib.import_module("orb-slam2-_b200")
B, W, H = 512, 640, 480
frames = pkg.pinned_empty((B, H, W), np.uint8)
base = [pkg.synth.frame(W, H, 1000 + i) for i in range(16)]
for i in range(B):
    frames[i] = np.roll(base[i % 16], i // 16, axis=1)
ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
cap = 1064
k, d, n = pkg.pinned_empty((B, cap), pkg.KP_DTYPE), pkg.pinned_empty((B, cap, 32), np.uint8), np.zeros(B, np.int32)
for i in range(4):
    print(f"--- call {i}", file=sys.stderr, flush=True)
    ex.extract_batch_into(frames, k, d, n)
