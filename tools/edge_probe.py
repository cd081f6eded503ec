import importlib, sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
pkg = importlib.import_module("orb-slam2-_b200")
import oracle_lib as ol
def run(name, img, nf=1000, nlev=8):
    try:
        ex = pkg.ORBextractor(nf,1.2,nlev,20,7,max_width=img.shape[1],max_height=img.shape[0])
        k,d = ex(img)
        try:
            ok,od = ol.Extractor(nf,1.2,nlev,20,7)(img)
            same = len(k)==len(ok) and np.array_equal(k.view(np.uint8),ok.view(np.uint8)) and np.array_equal(d,od)
        except Exception as e:
            same = f"oracle error {e}"
        print(name, "n", len(k), "== oracle:", same)
    except Exception as e:
        print(name, "ERROR", type(e).__name__, str(e)[:150])
run("black", np.zeros((480,640),np.uint8))
run("white", np.full((480,640),255,np.uint8))
run("noise_hi", np.random.default_rng(1).integers(0,256,(480,640),dtype=np.uint8))
run("checker", ((np.indices((480,640)).sum(0)//4)%2*255).astype(np.uint8))
run("small_160x120", pkg.synth.frame(160,120,3))
run("small_100x80", pkg.synth.frame(100,80,3))
run("small_64x48", pkg.synth.frame(64,48,3))
run("one_level", pkg.synth.frame(640,480,3), nlev=1)
run("few_features", pkg.synth.frame(640,480,3), nf=10)
run("many_features", pkg.synth.frame(640,480,3), nf=8000)
run("odd_641x479", pkg.synth.frame(641,479,3))
