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

w, h, seed, nf, nlev = 640, 480, 3, 1000, 1
img = pkg.synth.frame(w, h, seed)
ex = pkg.ORBextractor(nf, 1.2, nlev, 20, 7, max_width=w, max_height=h)
ox = oracle.Extractor(nf, 1.2, nlev, 20, 7)
k, d = ex(img); ok, od = ox(img)
for l in range(nlev):
    gc, oc = ex.level_candidates(l), ox.candidates(l)
    gs, os_ = set(map(tuple, gc.tolist())), set(map(tuple, oc.tolist()))
    print("level", l, len(gc), len(oc), "only gpu", sorted(gs - os_)[:20], "only oracle", sorted(os_ - gs)[:20])
    print("dups gpu", len(gc) - len(gs))
print("kp equal", np.array_equal(k.view(np.uint8), ok.view(np.uint8)), len(k), len(ok))
