"""Builds liborbx.so (the CUDA C-ABI library) in-tree with nvcc for sm_100a.

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box with the
repository snapshot. cudart is linked statically so the library depends on the driver only.
"""
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "liborbx.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# -fmad=false: nothing on the float paths may be contracted implicitly (bit-exact angles/descriptors)
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
         "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall", "-Xptxas", "-v"]


def sources():
    return sorted(CSRC.glob("*.cu"))


def _stale(target, deps):
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build(force=False, verbose=False):
    objdir = CSRC / "build"
    objdir.mkdir(exist_ok=True)
    headers = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list(CSRC.glob("*.inc")) + \
        [HERE.parent / "include" / "orbx.h", Path(__file__)]
    jobs = []
    for src in sources():
        obj = objdir / (src.stem + ".o")
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        r = subprocess.run([NVCC] + FLAGS + ["-c", str(src), "-o", str(obj)], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{r.stdout}\n{r.stderr}")
        (objdir / (src.stem + ".ptxas.txt")).write_text(r.stderr)
        return r.stderr

    with ThreadPoolExecutor(max_workers=8) as ex:
        logs = list(ex.map(compile_one, jobs))
    if verbose:
        for l in logs:
            print(l)
    objs = [objdir / (s.stem + ".o") for s in sources()]
    if force or jobs or _stale(LIB, objs):
        r = subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB)] +
                           [str(o) for o in objs], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    build_host_mirror(force)
    return LIB


def build_variant(name, defines):
    """Experiment aid (tools/variants.sh): the whole library again with extra -D flags ->
    orb-slam2-_b200/variants/liborbx_<name>.so, picked up by the Python mirror through ORBX_LIB=<path>."""
    objdir = CSRC / f"build_{name}"
    objdir.mkdir(exist_ok=True)
    out = HERE / "variants"
    out.mkdir(exist_ok=True)
    lib = out / f"liborbx_{name}.so"

    def compile_one(src):
        obj = objdir / (src.stem + ".o")
        r = subprocess.run([NVCC] + FLAGS + [f"-D{d}" for d in defines] + ["-c", str(src), "-o", str(obj)],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{r.stdout}\n{r.stderr}")
        (objdir / (src.stem + ".ptxas.txt")).write_text(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(compile_one, sources()))
    r = subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(lib)] +
                       [str(o) for o in objs], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return lib


def build_host_mirror(force=False):
    """host/libmirror_test.so: the C++ drop-in ORBextractor (host/ORBextractor.{h,cc}) compiled against the
    cv shim (OpenCV C++ is not installed here) plus C test entry points; linked to liborbx.so. Used by
    tests/test_gpu_host_mirror.py only -- a deployment compiles host/ORBextractor.cc against real OpenCV."""
    host = HERE / "host"
    out = host / "libmirror_test.so"
    srcs = [host / "ORBextractor.cc", host / "orbx_search.cc", host / "orbx_bow.cc", host / "mirror_capi.cc"]
    shim = HERE.parent / "oracle" / "ref_shim"
    deps = srcs + [host / "ORBextractor.h", host / "orbx_search.h", host / "orbx_bow.h", host / "orbx_shard.h", HERE.parent / "include" / "orbx.h", shim / "opencv2" / "core" / "core.hpp", LIB]
    if not (force or _stale(out, deps)):
        return out
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++11", "-fPIC", "-shared", "-Wall", "-Wno-stringop-overflow", "-Wno-array-bounds",
           f"-I{host}", f"-I{HERE.parent / 'include'}", f"-I{shim}"] + [str(s) for s in srcs] + \
          ["-o", str(out), f"-L{HERE}", "-lorbx", "-Wl,-rpath,$ORIGIN/.."]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"host mirror build failed:\n{r.stdout}\n{r.stderr}")
    return out


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
