#!/usr/bin/env python3
"""Streaming host-buffer path (orbx_extract_batch_submit / _wait): frames/s vs chunk size x lanes x steps in flight,
next to the synchronous orbx_extract_batch call. 640x480, 1000 features, 512 frames per step, pinned buffers."""
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import orbslam2_b200 as pkg  # noqa: E402

W, H, B, NF, STEPS = 640, 480, 512, 1000, 12
base = np.stack([pkg.synth.frame(W, H, i) for i in range(32)])
frames = pkg.pinned_empty((B, H, W), np.uint8)
for i in range(B):
    frames[i] = np.roll(base[i % 32], i // 32, 1)
cap = NF + 64
sets = [(pkg.pinned_empty((B, cap), pkg.KP_DTYPE), pkg.pinned_empty((B, cap, 32), np.uint8), pkg.pinned_empty(B, np.int32))
        for _ in range(4)]
ex = pkg.ORBextractor(NF, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)


def sync_fps():
    for _ in range(3):
        ex.extract_batch_into(frames, *sets[0])
    t0 = time.perf_counter()
    for _ in range(STEPS):
        ex.extract_batch_into(frames, *sets[0])
    return B * STEPS / (time.perf_counter() - t0)


def stream_fps(depth):
    def run(k):
        pending = [None] * depth
        for i in range(k):
            if pending[i % depth] is not None:
                ex.extract_batch_wait(pending[i % depth])
            pending[i % depth] = ex.extract_batch_submit(frames, *sets[i % depth])
        for t in pending:
            if t is not None:
                ex.extract_batch_wait(t)
    run(3)
    t0 = time.perf_counter()
    run(STEPS)
    return B * STEPS / (time.perf_counter() - t0)


for chunk, lanes in ((64, 4), (64, 3), (128, 4), (128, 3), (128, 2), (256, 2), (256, 3), (32, 4)):
    ex.set_pipeline(chunk_frames=chunk, lanes=lanes)
    line = f"chunk {chunk:3d} lanes {lanes}: sync {sync_fps():9.0f} fps"
    for depth in (2, 3):
        line += f" | stream depth {depth} {stream_fps(depth):9.0f} fps"
    print(line, flush=True)
