"""Pinned host<->device copy bandwidth of this box (context for the e2e number: 640x480 frames are 307 KB each)."""
import time
import torch
n = 256 * 307200
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, f in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        f()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    print(f"{name}: {n / dt / 1e9:.1f} GB/s  ({n / 307200 / dt:.0f} frames/s of 640x480)")
