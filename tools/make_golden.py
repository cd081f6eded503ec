#!/usr/bin/env python3
"""Generate tests/golden/*.npz in the authoring container.

prims_cv2.npz : inputs and the outputs of the REAL OpenCV (cv2 4.13.0, the only OpenCV on either
                box) for the five primitives the reference calls: cv::resize INTER_LINEAR, cv::GaussianBlur
                7x7 s=2 REFLECT_101, cv::FAST 9/16 with NMS at thresholds 20 and 7, cv::fastAtan2.
                These pin the oracle's restatements (tests/test_golden.py) without needing cv2 at test time.
extract_oracle.npz : sha256 of the oracle's own extractor output on seeded synthetic frames (regression
                pin of the canonical-order extractor; NOT an independent reference).
"""
import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import cv2  # noqa: E402
import oracle_lib as O  # noqa: E402
import orbslam2_b200 as pkg  # noqa: E402  (synth only)

OUT = ROOT / "tests" / "golden"


def fast_cv2(img, th):
    det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True,
                                         type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    kps = det.detect(np.ascontiguousarray(img))
    return np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], np.int32).reshape(-1, 3)


def main():
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    cv2.setNumThreads(1)
    OUT.mkdir(parents=True, exist_ok=True)
    d = {"cv2_version": np.array(cv2.__version__)}
    imgs = {"shapes": pkg.synth.frame(120, 90, 101), "noise": pkg.synth.frame(97, 61, 102, "noise")}
    for name, im in imgs.items():
        d[f"img_{name}"] = im
        h, w = im.shape
        for s in (1.2, 1.728):
            dw, dh = int(round(w / s)), int(round(h / s))
            d[f"resize_{name}_{dw}x{dh}"] = cv2.resize(im, (dw, dh), interpolation=cv2.INTER_LINEAR)
        d[f"blur_{name}"] = cv2.GaussianBlur(im, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        for th in (20, 7):
            d[f"fast_{name}_{th}"] = fast_cv2(im, th)
        # the per-cell call shape of the reference: a 37x38 ROI with a row stride larger than its width
        d[f"fastcell_{name}_7"] = fast_cv2(im[11:49, 20:57], 7)
    rng = np.random.default_rng(7)
    y = rng.integers(-60000, 60000, 4096).astype(np.float32)
    x = rng.integers(-60000, 60000, 4096).astype(np.float32)
    y[:6] = [0, 0, 5, -5, 1, -1]
    x[:6] = [0, 5, 0, 0, 1, -1]
    d["atan2_y"], d["atan2_x"] = y, x
    d["atan2_deg"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    np.savez_compressed(OUT / "prims_cv2.npz", **d)

    e = {}
    for (w, h, nf, seed, kind) in [(640, 480, 1000, 0, "shapes"), (640, 480, 1000, 1, "shapes"),
                                   (752, 480, 1200, 2, "shapes"), (1241, 376, 2000, 3, "shapes"),
                                   (640, 480, 1000, 4, "noise"), (640, 480, 1000, 5, "lowcontrast")]:
        ex = O.Extractor(nf, 1.2, 8, 20, 7)
        k, dd = ex(pkg.synth.frame(w, h, seed, kind))
        key = f"{w}x{h}_{nf}_{seed}_{kind}"
        e[key + "_n"] = np.array(len(k))
        e[key + "_sha"] = np.array(hashlib.sha256(k.tobytes() + dd.tobytes()).hexdigest())
        e[key + "_imgsha"] = np.array(hashlib.sha256(pkg.synth.frame(w, h, seed, kind).tobytes()).hexdigest())
    np.savez_compressed(OUT / "extract_oracle.npz", **e)
    print("wrote", [p.name for p in OUT.iterdir()])


if __name__ == "__main__":
    main()
