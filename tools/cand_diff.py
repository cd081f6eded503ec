"""Debug aid: candidate sets of the CUDA FAST kernel vs the oracle for one frame / configuration."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import importlib.util
root = os.path.join(os.path.dirname(__file__), "..")
spec = importlib.util.spec_from_file_location("orbslam2_b200", os.path.join(root, "orb-slam2-_b200", "__init__.py"),
                                              submodule_search_locations=[os.path.join(root, "orb-slam2-_b200")])
pkg = importlib.util.module_from_spec(spec); sys.modules["orbslam2_b200"] = pkg; spec.loader.exec_module(pkg)
import oracle_lib as oracle

# usage: cand_diff.py [w h seed nf nlev scale ini min kind]
a = sys.argv[1:]
w, h, seed, nf, nlev = (int(v) for v in a[:5]) if len(a) >= 5 else (640, 480, 3, 1000, 1)
scale = float(a[5]) if len(a) > 5 else 1.2
ini, mn = (int(a[6]), int(a[7])) if len(a) > 7 else (20, 7)
kind = a[8] if len(a) > 8 else "shapes"
img = pkg.synth.frame(w, h, seed, kind=kind)
ex = pkg.ORBextractor(nf, scale, nlev, ini, mn, max_width=w, max_height=h)
ox = oracle.Extractor(nf, scale, nlev, ini, mn)
k, d = ex(img); ok, od = ox(img)
for l in range(nlev):
    gc, oc = ex.level_candidates(l), ox.candidates(l)
    gs, os_ = set(map(tuple, gc.tolist())), set(map(tuple, oc.tolist()))
    print("level", l, len(gc), len(oc), "only gpu", sorted(gs - os_)[:20], "only oracle", sorted(os_ - gs)[:20])
    print("dups gpu", len(gc) - len(gs))
    miss = np.array(sorted(os_ - gs)).reshape(-1, 3)
    if len(miss):
        xs, cnt = np.unique(miss[:, 0], return_counts=True)
        print("missing by x:", dict(zip(xs.tolist(), cnt.tolist())))
        ys, cnt = np.unique(miss[:, 1] // 8, return_counts=True)
        print("missing by y//8:", dict(zip(ys.tolist(), cnt.tolist())))
        oc_a = np.array(sorted(os_)).reshape(-1, 3)
        for x in xs.tolist()[:3]:
            print("oracle candidates at x =", x, ":", int((oc_a[:, 0] == x).sum()), "gpu:", int((np.array(sorted(gs)).reshape(-1, 3)[:, 0] == x).sum()))
print("kp equal", np.array_equal(k.view(np.uint8), ok.view(np.uint8)), len(k), len(ok))
