"""Latency of the single-frame entry points (what Frame::Frame calls once per image): orbx_extract on pageable and
pinned frames, the stereo pair on two handles from two threads, and the ctypes overhead."""
import importlib
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
pkg = importlib.import_module("orb-slam2-_b200")
from importlib import import_module
capi = import_module("orb-slam2-_b200._capi")


def timeit(fn, reps=200):
    for _ in range(20):
        fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps * 1e6


for (w, h, nf) in ((640, 480, 1000), (752, 480, 1200), (1241, 376, 2000)):
    img = pkg.synth.frame(w, h, 3)
    pimg = pkg.pinned_empty((h, w), np.uint8)
    pimg[:] = img
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    cap = nf + 64
    k, d, n = pkg.pinned_empty((1, cap), pkg.KP_DTYPE), pkg.pinned_empty((1, cap, 32), np.uint8), np.zeros(1, np.int32)
    pk, pd = np.empty((1, cap), pkg.KP_DTYPE), np.empty((1, cap, 32), np.uint8)
    t_page = timeit(lambda: ex.extract_batch_into(img[None], pk, pd, n))
    t_pin = timeit(lambda: ex.extract_batch_into(pimg[None], k, d, n))
    ex.set_profiling(True)
    ex.extract_batch_into(pimg[None], k, d, n)
    dev = ex.last_stage_ms()["total"] * 1e3
    ex.set_profiling(False)
    print(f"{w}x{h}/{nf}: orbx_extract pageable {t_page:7.1f} us   pinned in/out {t_pin:7.1f} us   kernels {dev:6.1f} us", flush=True)
    # stereo: two handles, two host threads, like R/src/Frame.cpp:86-89
    ex2 = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    k2, d2, n2 = pkg.pinned_empty((1, cap), pkg.KP_DTYPE), pkg.pinned_empty((1, cap, 32), np.uint8), np.zeros(1, np.int32)

    def pair():
        th = threading.Thread(target=lambda: ex2.extract_batch_into(pimg[None], k2, d2, n2))
        th.start()
        ex.extract_batch_into(pimg[None], k, d, n)
        th.join()
    print(f"    stereo pair on two handles / two threads: {timeit(pair, 100):7.1f} us per pair", flush=True)
