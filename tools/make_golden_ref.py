#!/usr/bin/env python3
"""Generate tests/golden/ref_extract.npz from the REFERENCE's own ORBextractor.cpp (oracle/_ref, monotonic
allocator = canonical octree tie-break). Run in the authoring container, where /root/reference exists:

    make -C oracle && python tools/make_golden_ref.py

Stores, per seeded synthetic frame, the keypoint count, a sha256 over keypoints+descriptors, the first 8
keypoints/descriptors verbatim and a sha256 of the frame (guards the synthetic generator). The GPU box has
no /root/reference: there these vectors are what ties the CUDA path to the reference.
"""
import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import ref_lib  # noqa: E402
import orbslam2_b200 as pkg  # noqa: E402  (synth only)

CASES = [(640, 480, 1000, 0, "shapes"), (640, 480, 1000, 1, "shapes"), (752, 480, 1200, 2, "shapes"),
         (1241, 376, 2000, 3, "shapes"), (640, 480, 1000, 4, "noise"), (640, 480, 1000, 5, "lowcontrast"),
         (1920, 1080, 4000, 6, "shapes")]


def main():
    assert ref_lib.available(), "build oracle/_ref first (make -C oracle)"
    ref_lib.set_allocator_mode(1)
    e = {}
    for (w, h, nf, seed, kind) in CASES:
        img = pkg.synth.frame(w, h, seed, kind)
        k, d = ref_lib.RefExtractor(nf, 1.2, 8, 20, 7)(img)
        key = f"{w}x{h}_{nf}_{seed}_{kind}"
        e[key + "_n"] = np.array(len(k))
        e[key + "_sha"] = np.array(hashlib.sha256(k.tobytes() + d.tobytes()).hexdigest())
        e[key + "_imgsha"] = np.array(hashlib.sha256(img.tobytes()).hexdigest())
        e[key + "_k8"] = k[:8].copy()
        e[key + "_d8"] = d[:8].copy()
        print(key, len(k))
    np.savez_compressed(ROOT / "tests" / "golden" / "ref_extract.npz", **e)


if __name__ == "__main__":
    main()
