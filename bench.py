#!/usr/bin/env python3
"""bench.py -- the driver's measurement contract for the ORB front-end hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Top level of the JSON line = BASELINE.json `metric`, config C1: ORBextractor::operator() on 640x480 u8 frames,
1000 features, 8 levels, scale 1.2, FAST 20/7 (TUM1.yaml). One "step" = one pass of the extraction pipeline over a
batch of distinct synthetic frames larger than the 126 MB L2. Frames shard across ranks with no data-path collective
("weak" scaling: the per-GPU batch is fixed).

  value        : frames/s, inputs and outputs resident in HBM (orbx_extract_batch_device)
  e2e          : frames/s through the host-buffer C ABI (orbx_extract_batch_submit/_wait: pinned host frames -> H2D ->
                 kernels -> D2H keypoints + descriptors), with the PCIe ceiling measured in the same run at the same N
  roofline / stages : per-stage CUDA-event times from inside the timed region
  parity       : sampled frames of the timed batch compared bit for bit with the CPU oracle (outside the timed region)
  cpu_baseline : the reference's own ORBextractor.cpp (oracle/_ref) on this box's host cores (rank 0, N = 1 only), plus
                 the same primitives through cv2 4.13 on one thread
  workloads    : the other extraction configs of BASELINE.json at the same N -- C2 (KITTI stereo pairs + row matching),
                 C3 (EuRoC 752x480 stereo SEQUENCE of 10 000 frames streamed from pinned host memory, sharded pair-aligned
                 across the ranks), C4 (1920x1080 / 4000 features) -- each with value, e2e, stages, roofline, parity
  hamming      : config 5, 10k x 10M brute-force top-2 + ratio through orbx_shard_match_bf2_device (database sharded
                 across the ranks, ncclAllGather issued by liborbx.so, merge kernel), with parity, roofline and CPU baseline

--impl reference times the reference's own CPU code (oracle/_ref/libref_orbextractor.so: R/src/ORBextractor.cpp compiled
against scalar restatements of the OpenCV primitives; the oracle port when that library is absent) with all host threads
on the same config. Outside tests/ and smoke() this file is the only place where oracle/ is executed, and only as the
checker (parity blocks) or as the thing timed on the CPU (cpu_baseline legs, --impl reference).
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

NLEVELS, SCALE, INI_TH, MIN_TH = 8, 1.2, 20, 7
BASE_FRAMES = 32               # distinct numpy-generated frames; the rest are shifted / noised copies

# BASELINE.json configs 1-4. frames = frames per step per GPU (about 150 MB of distinct input, > the 126 MB L2).
WORKLOADS = {
    "C1": dict(w=640, h=480, nfeat=1000, frames=512, stereo=False, mbf=0.0,
               metric="ORB frames/s (640x480, 1000 kp)",
               workload="C1: ORBextractor::operator() on 640x480 u8 frames, 1000 features, 8 levels, scale 1.2, "
                        "FAST 20/7 (TUM1.yaml)"),
    "C2": dict(w=1241, h=376, nfeat=2000, frames=320, stereo=True, mbf=float(np.float32(718.856) * np.float32(0.537)),
               metric="ORB frames/s (1241x376 stereo pairs, 2000 kp)",
               workload="C2: KITTI 1241x376 stereo pairs (L R L R ...), 2000 features per image, both extractions + "
                        "Frame::ComputeStereoMatches per pair (KITTI00-02.yaml)"),
    "C3": dict(w=752, h=480, nfeat=1200, frames=416, stereo=True, mbf=float(np.float32(435.2) * np.float32(0.11)),
               sequence=10_000,
               metric="ORB frames/s (752x480 stereo sequence, 1200 kp)",
               workload="C3: EuRoC 752x480 stereo sequence of 10000 frames (5000 L/R pairs), 1200 features, sharded "
                        "pair-aligned across the ranks (EuRoC.yaml)"),
    "C4": dict(w=1920, h=1080, nfeat=4000, frames=72, stereo=False, mbf=0.0,
               metric="ORB frames/s (1920x1080, 4000 kp)",
               workload="C4: 1920x1080 frames, 4000 features, 8 levels, batched"),
}
HAM_NQ, HAM_NDB = 10_000, 10_000_000   # BASELINE.json config 5
HAM_BLOCK = 125_000                    # database rows are generated in seeded blocks: any rank can rebuild any block


def static_config(cfg):
    """The `config` object: identical in both arms (the driver compares them)."""
    return {"workload": cfg["workload"], "shape": f'{cfg["w"]}x{cfg["h"]}', "nfeatures": cfg["nfeat"], "nlevels": NLEVELS,
            "scale_factor": SCALE, "fast": f"{INI_TH}/{MIN_TH}", "frames_per_step_per_gpu": cfg["frames"],
            "l2_policy": "inputs larger than L2 (about 150 MB of distinct frames per step)",
            "parallelism": "frame sharding across the GPUs, no collective"}


def level_sizes(w, h):
    sc = [np.float32(1.0)]
    for _ in range(1, NLEVELS):
        sc.append(np.float32(np.float64(sc[-1]) * np.float64(np.float32(SCALE))))
    inv = [np.float32(1.0) / s for s in sc]
    return [(int(np.rint(np.float32(w) * s)), int(np.rint(np.float32(h) * s))) for s in inv]


def algorithmic_bytes(cfg):
    """bytes per frame, per stage and total (5 702 474 B at C1)."""
    A = [w * h for w, h in level_sizes(cfg["w"], cfg["h"])]
    sA = sum(A)
    st = {"pyramid": (sA - A[-1]) + (sA - A[0]), "fast": sA, "blur": 2 * sA,
          "describe": cfg["nfeat"] * (709 + 512) + cfg["nfeat"] * 60, "octree": 0}
    st["total"] = st["pyramid"] + st["fast"] + st["blur"] + st["describe"]
    return st


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def base_frames(pkg, cfg, n=BASE_FRAMES, seed0=1000):
    w, h = cfg["w"], cfg["h"]
    if w * h > 1_000_000:
        n = min(n, 12)
    if cfg["stereo"]:   # stereo pairs, interleaved L R L R ...
        return np.stack([im for i in range(n // 2) for im in pkg.synth.stereo_pair(w, h, seed0 + i)])
    return np.stack([pkg.synth.frame(w, h, seed0 + i) for i in range(n)])


def device_frames(torch, base, first, count, salt=0):
    """Frames first .. first+count-1 of an endless sequence of distinct frames on the device: frame i is base frame
    i % nb rolled by an offset that grows with i // nb, plus +-1 noise from a generator seeded by (salt, i // nb) --
    a function of the frame index only, so that every sharding of a sequence sees the same frames. A stereo pair
    (two consecutive base frames) gets the same roll on both sides, which keeps its disparities. Generation happens
    outside every timed region."""
    dev = torch.device("cuda", torch.cuda.current_device())
    b = torch.from_numpy(base).to(dev)
    nb = b.shape[0]
    out = torch.empty((count,) + tuple(b.shape[1:]), dtype=torch.uint8, device=dev)
    g = torch.Generator(device=dev)
    i = first
    while i < first + count:
        rep, j = divmod(i, nb)
        m = min(nb - j, first + count - i)
        fr = torch.roll(b[j:j + m], shifts=(3 * rep, 5 * rep), dims=(1, 2)).to(torch.int16)
        if rep:
            g.manual_seed(1234 + 7919 * salt + rep)
            noise = torch.randint(-1, 2, (nb,) + tuple(b.shape[1:]), generator=g, device=dev, dtype=torch.int16)
            fr = fr + noise[j:j + m]
        out[i - first:i - first + m] = fr.clamp_(0, 255).to(torch.uint8)
        i += m
    return out


# ---------------------------------------------------------------------------------------------------------------------
# CPU legs (the only code here that runs oracle/)
# ---------------------------------------------------------------------------------------------------------------------
def cpu_extractor_factory(cfg):
    """(constructor, kind): the reference's own ORBextractor.cpp (oracle/_ref, built in the authoring container from
    /root/reference against the cv shim; the .so travels) when present -> kind "reference"; else the oracle port."""
    import ref_lib
    nf = cfg["nfeat"]
    if ref_lib.available():
        ref_lib.lib()
        return (lambda: ref_lib.RefExtractor(nf, SCALE, NLEVELS, INI_TH, MIN_TH)), "reference"
    import oracle_lib
    oracle_lib.lib()
    return (lambda: oracle_lib.Extractor(nf, SCALE, NLEVELS, INI_TH, MIN_TH)), "port"


CPU_NOTE = ("kind 'reference' = the reference's own ORBextractor.cpp control flow compiled where it lies, linked against "
            "SCALAR restatements of cv::FAST / resize / GaussianBlur (oracle/cv_prims.cpp; OpenCV C++ cannot be installed "
            "here). A real OpenCV 4.13 build runs those primitives with SIMD: see opencv_primitives_ms -- the reference arm "
            "is therefore pessimistic by about 2x and every 'x CPU' ratio should be read with that factor")


def time_cpu_extractor(cfg, frames, budget_s, nthreads):
    """frames/s of the CPU reference with `nthreads` threads, each with its own extractor instance
    (the reference's own pattern: one ORBextractor per std::thread, R/src/Frame.cpp:86-89)."""
    nfr = len(frames)
    make = cpu_extractor_factory(cfg)[0]
    exs = [make() for _ in range(nthreads)]
    exs[0](frames[0])  # warm caches / page in
    done = [0] * nthreads
    per_call = []
    stop = time.perf_counter() + budget_s

    def work(t):
        i = t
        while time.perf_counter() < stop:
            t0 = time.perf_counter()
            exs[t](frames[i % nfr])
            if t == 0:
                per_call.append(time.perf_counter() - t0)
            done[t] += 1
            i += nthreads

    t0 = time.perf_counter()
    ths = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    dt = time.perf_counter() - t0
    return sum(done) / dt, sum(done), float(np.median(per_call)) * 1e3 if per_call else None


def opencv_primitives_ms(cfg, frames):
    """The three OpenCV primitives that dominate the reference's CPU time, through the real (SIMD) cv2 4.13 on ONE
    thread, on the same frames: 7 resizes, per-level FAST at iniTh and again at minTh (the reference runs minTh only in
    the cells that stay empty, so this is an upper bound of its FAST time), 8 clone + GaussianBlur. ms per frame."""
    try:
        import cv2
    except ImportError:
        return None
    cv2.setNumThreads(1)
    sizes = level_sizes(cfg["w"], cfg["h"])
    f_ini = cv2.FastFeatureDetector_create(INI_TH, True)
    f_min = cv2.FastFeatureDetector_create(MIN_TH, True)
    t = {"resize": 0.0, "fast_ini": 0.0, "fast_min": 0.0, "blur": 0.0}
    reps = 0
    t_end = time.perf_counter() + 3.0
    while time.perf_counter() < t_end or reps < 4:
        img = frames[reps % len(frames)]
        t0 = time.perf_counter()
        pyr = [img]
        for l in range(1, NLEVELS):
            pyr.append(cv2.resize(pyr[-1], sizes[l], interpolation=cv2.INTER_LINEAR))
        t1 = time.perf_counter()
        for p in pyr:
            f_ini.detect(p[16:-16, 16:-16])
        t2 = time.perf_counter()
        for p in pyr:
            f_min.detect(p[16:-16, 16:-16])
        t3 = time.perf_counter()
        for p in pyr:
            cv2.GaussianBlur(p.copy(), (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        t4 = time.perf_counter()
        t["resize"] += t1 - t0
        t["fast_ini"] += t2 - t1
        t["fast_min"] += t3 - t2
        t["blur"] += t4 - t3
        reps += 1
    out = {k: v / reps * 1e3 for k, v in t.items()}
    out["total_ini_only"] = out["resize"] + out["fast_ini"] + out["blur"]
    out["total_with_min_pass"] = out["total_ini_only"] + out["fast_min"]
    out["what"] = f"cv2 {cv2.__version__}, setNumThreads(1), {reps} frames: ms per frame"
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on all host threads, same config / metric."""
    import orbslam2_b200 as pkg  # synth only; no CUDA call is made on this path
    cfg = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    frames = [f for f in base_frames(pkg, cfg)]
    budget = 4.0
    for _ in range(min(args.warmup, 1)):
        time_cpu_extractor(cfg, frames, 1.0, cores)
    t0 = time.perf_counter()
    tot_frames, vals, med_ms = 0, [], None
    for _ in range(args.steps):
        fps, n, med_ms = time_cpu_extractor(cfg, frames, budget, cores)
        vals.append(fps)
        tot_frames += n
    dt = time.perf_counter() - t0
    value = float(np.mean(vals))
    sample = (f"{args.steps} steps x {budget:.0f} s of {cfg['w']}x{cfg['h']} frames on {cores} threads ({tot_frames} frames), "
              "one extractor per thread")
    line = {"impl": "reference", "metric": cfg["metric"], "value": value, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": static_config(cfg),
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": cpu_extractor_factory(cfg)[1],
                             "sample": sample, "single_thread_ms_per_frame": med_ms, "note": CPU_NOTE,
                             "opencv_primitives_ms": opencv_primitives_ms(cfg, frames)},
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def hamming_cpu_baseline(pkg, q, db_slice):
    """the reference's DescriptorDistance inside the running best / second-best loop
    (oracle port of R/src/ORBmatcher.cpp:1901-1917, :266-300) on a slice of config 5, one thread and all cores."""
    import oracle_lib as ol
    cores = os.cpu_count() or 1
    rows = db_slice.shape[0]
    q1 = q[:500]
    t0 = time.perf_counter()
    ol.match_bf2(q1, db_slice, np.float32(0.6), 50, nthreads=1)
    t1 = time.perf_counter() - t0
    t0 = time.perf_counter()
    ol.match_bf2(q, db_slice, np.float32(0.6), 50, nthreads=cores)
    tn = time.perf_counter() - t0
    g1 = len(q1) * rows / t1 / 1e9
    gn = len(q) * rows / tn / 1e9
    return {"value": gn, "unit": "Gpairs/s", "cores": cores, "kind": "port",
            "single_thread_gpairs_s": g1,
            "sample": f"{len(q)} queries x {rows} rows on {cores} threads ({tn:.1f} s), {len(q1)} x {rows} on one thread "
                      f"({t1:.1f} s); the full {HAM_NQ} x {HAM_NDB} problem is this x{HAM_NDB // rows} "
                      f"(extrapolated: {HAM_NQ * HAM_NDB / (gn * 1e9):.0f} s on all cores)"}


# ---------------------------------------------------------------------------------------------------------------------
# GPU legs
# ---------------------------------------------------------------------------------------------------------------------
class Ctx:
    pass


def make_ctx(torch):
    import torch.distributed as dist
    c = Ctx()
    c.torch, c.dist = torch, dist
    c.world = int(os.environ.get("WORLD_SIZE", "1"))
    c.rank = int(os.environ.get("RANK", "0"))
    c.local = int(os.environ.get("LOCAL_RANK", "0"))
    if c.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", c.local))

    def barrier():
        if c.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if c.world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    c.barrier = barrier
    c.max_over_ranks = lambda x: reduce(x, dist.ReduceOp.MAX)
    c.sum_over_ranks = lambda x: reduce(x, dist.ReduceOp.SUM)
    c.min_over_ranks = lambda x: reduce(x, dist.ReduceOp.MIN)
    c.stream = torch.cuda.current_stream().cuda_stream
    assert c.stream != 0
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    if peaks_path.exists():
        c.peak, c.peak_src = json.loads(peaks_path.read_text())["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    else:
        c.peak, c.peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    return c


def pcie_probe(c, h2d_bytes, d2h_bytes, iters=8):
    """What this box's PCIe / host memory delivers with all `world` ranks copying at once: pinned H2D pieces of the
    step's input size and D2H pieces of its result size on two streams, both directions concurrently (the pattern of
    the streaming pipeline), timed with CUDA events on each stream. Aggregate GB/s over the ranks."""
    torch = c.torch
    src = torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory()
    dst = torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(h2d_bytes, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(d2h_bytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(both):
        c.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        with torch.cuda.stream(s1):
            ev[0].record()
            for _ in range(iters):
                d_in.copy_(src, non_blocking=True)
            ev[1].record()
        if both:
            with torch.cuda.stream(s2):
                ev[2].record()
                for _ in range(iters):
                    dst.copy_(d_out, non_blocking=True)
                ev[3].record()
        c.barrier()
        h = h2d_bytes * iters / (ev[0].elapsed_time(ev[1]) * 1e-3) / 1e9
        d = d2h_bytes * iters / (ev[2].elapsed_time(ev[3]) * 1e-3) / 1e9 if both else 0.0
        return h, d

    run(True)
    h_only, _ = run(False)
    h_both, d_both = run(True)
    return {"h2d_gbs_alone": c.sum_over_ranks(h_only), "h2d_gbs_with_d2h": c.sum_over_ranks(h_both),
            "d2h_gbs_with_h2d": c.sum_over_ranks(d_both), "per_rank_h2d_gbs_min": c.min_over_ranks(h_both),
            "piece_bytes": [h2d_bytes, d2h_bytes], "ranks_copying_at_once": c.world}


def kept_rows_per_frame(ex, cfg):
    """rows the device keeps per frame (sum over the levels of keptCap = max(quota + 3, 4 * nIni, 64) + 1):
    what one frame's D2H copy moves"""
    nini = int(np.rint(cfg["w"] / cfg["h"]))
    return sum(max(f + 3, 4 * nini, 64) + 1 for f in ex.mnFeaturesPerLevel.tolist())


def unpack_kps(pkg, a):
    """(B, cap, 7) float32 device tensor -> (B, cap) KP_DTYPE host array"""
    b, cap = a.shape[:2]
    return a.cpu().numpy().view(np.uint8).reshape(b, cap, 28).view(pkg.KP_DTYPE).reshape(b, cap)


def oracle_check(pkg, cfg, frames, got, pairs=None):
    """Bit-for-bit comparison of sampled frames with the CPU oracle. frames: {index: (h, w) uint8}; got[index] =
    (kps, desc) of the path under test. pairs: {(iL, iR): (u_right, depth, n_matched)} -> Frame::ComputeStereoMatches."""
    import oracle_lib as ol
    ox = ol.Extractor(cfg["nfeat"], SCALE, NLEVELS, INI_TH, MIN_TH)
    res = {}
    bad = []
    for i, img in frames.items():
        k, d = ox(img)
        pyr = [ox.level(l).copy() for l in range(NLEVELS)] if pairs else None
        res[i] = (k, d, pyr)
        gk, gd = got[i]
        if len(gk) != len(k) or not np.array_equal(gk.view(np.uint8), k.view(np.uint8)) or not np.array_equal(gd, d):
            bad.append(int(i))
    npairs = 0
    if pairs:
        t = ox.tables()
        for (il, ir), (u, z, nm) in pairs.items():
            kl, dl, pl = res[il]
            kr, dr, pr = res[ir]
            eu, ez, en = ol.stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], np.float32(cfg["mbf"]), np.float32(0.0))
            npairs += 1
            if en != nm or not np.array_equal(u[:len(kl)].view(np.uint32), eu.view(np.uint32)) or \
                    not np.array_equal(z[:len(kl)].view(np.uint32), ez.view(np.uint32)):
                bad.append([int(il), int(ir)])
    return {"frames_checked": len(frames), "stereo_pairs_checked": npairs, "mismatches": bad}


def frame_digests(n, kps, desc, ur=None, dp=None, nm=None):
    """8-byte digest per frame of everything the path returned for it (count, keypoint rows, descriptor rows and, for
    the left frame of a pair, mvuRight / mvDepth rows + the match count)."""
    out = np.empty(len(n), np.uint64)
    for f in range(len(n)):
        h = hashlib.blake2b(digest_size=8)
        nf = int(n[f])
        h.update(int(nf).to_bytes(4, "little", signed=True))
        h.update(kps[f, :max(nf, 0)].tobytes())
        h.update(desc[f, :max(nf, 0)].tobytes())
        if ur is not None and f % 2 == 0:
            h.update(ur[f // 2, :max(nf, 0)].tobytes())
            h.update(dp[f // 2, :max(nf, 0)].tobytes())
            h.update(int(nm[f // 2]).to_bytes(4, "little", signed=True))
        out[f] = int.from_bytes(h.digest(), "little")
    return out


def run_extraction(c, pkg, name, args, steps, with_sync_call=False):
    """One extraction config at the current N: device-resident value, end-to-end value through the host-buffer C ABI,
    stage times, roofline, parity against the oracle."""
    torch = c.torch
    cfg = WORKLOADS[name]
    W, H, NFEAT, B, stereo = cfg["w"], cfg["h"], cfg["nfeat"], cfg["frames"], cfg["stereo"]
    ex = pkg.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=c.local, max_width=W, max_height=H, max_batch=B)
    base = base_frames(pkg, cfg)
    frames = device_frames(torch, base, 0, B, salt=c.rank + 1)     # every rank works on frames of its own
    cap = NFEAT + 64
    d_kps = torch.empty((B, cap, 7), dtype=torch.float32, device="cuda")
    d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
    stream = c.stream
    pairs = B // 2
    if stereo:
        # the batch is a sequence of stereo pairs (L R L R ...); Frame::ComputeStereoMatches for every pair runs on the
        # same stream right after the extraction (mb = 0: what the reference's stereo constructor runs with)
        sm = pkg.ORBmatcher(device=c.local)
        d_ur = torch.empty((pairs, cap), dtype=torch.float32, device="cuda")
        d_z = torch.empty((pairs, cap), dtype=torch.float32, device="cuda")
        d_nm = torch.zeros(pairs, dtype=torch.int32, device="cuda")

    def step():
        ex.extract_batch_device(frames.data_ptr(), B, W, H, W, W * H, d_kps.data_ptr(), d_desc.data_ptr(), cap,
                                d_n.data_ptr(), stream)
        if stereo:
            sm.stereo_match_batch_device(ex, pairs, 0, 1, 2, d_kps.data_ptr(), d_desc.data_ptr(), d_n.data_ptr(), cap,
                                         cfg["mbf"], 0.0, d_ur.data_ptr(), d_z.data_ptr(), d_nm.data_ptr(), stream)

    # ---- device-resident throughput ("value") ----
    for _ in range(max(args.warmup, 3)):
        step()
    ex.set_profiling(True)
    stage_acc = {}
    sampler = ClockSampler(c.local)
    c.barrier()
    if c.rank == 0:
        sampler.start()
    launches0 = ex.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
        for k, v in ex.last_stage_ms().items():  # CUDA events recorded on the launching stream
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    e1.record()
    c.barrier()
    ms = c.max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if c.rank == 0 else None
    launches = ex.launch_count() - launches0 + (3 * steps if stereo else 0)
    ex.set_profiling(False)
    value = c.world * B * steps / (ms * 1e-3)
    n_dev = d_n.cpu().numpy()
    overflowed = int((n_dev < 0).sum())     # orbx_extract_batch_device marks truncated frames with n = -1
    nkp = int(n_dev.clip(min=0).sum())

    # ---- end to end through the host-buffer C ABI ("e2e"): the frames as a continuous sequence through
    # orbx_extract_batch_submit / orbx_stereo_batch_submit + _wait, two steps in flight, each step into its own pinned
    # result arrays; a step's results are waited for before its arrays are handed out again
    e2e_obj, h_sets, hnp = None, None, None
    if not args.skip_e2e:
        host = torch.empty((B, H, W), dtype=torch.uint8).pin_memory()
        host.copy_(frames.cpu())
        hnp = host.numpy()

        def result_set():
            s = [pkg.pinned_empty((B, cap), pkg.KP_DTYPE, device=c.local), pkg.pinned_empty((B, cap, 32), np.uint8, device=c.local),
                 pkg.pinned_empty(B, np.int32, device=c.local)]
            if stereo:
                s += [pkg.pinned_empty((pairs, cap), np.float32, device=c.local), pkg.pinned_empty((pairs, cap), np.float32, device=c.local),
                      pkg.pinned_empty(pairs, np.int32, device=c.local)]
            return s
        h_sets = [result_set(), result_set()]

        def submit(s):
            if stereo:
                return ex.stereo_batch_submit(hnp, s[0], s[1], s[2], cfg["mbf"], 0.0, s[3], s[4], s[5])
            return ex.extract_batch_submit(hnp, s[0], s[1], s[2])

        def run_stream(k):
            pending = [None, None]
            for i in range(k):
                if pending[i % 2] is not None:
                    ex.extract_batch_wait(pending[i % 2])
                pending[i % 2] = submit(h_sets[i % 2])
            for t in pending:
                if t is not None:
                    ex.extract_batch_wait(t)
        l0 = ex.launch_count()
        run_stream(3)
        per_step_launches = (ex.launch_count() - l0) // 3
        c.barrier()
        t0 = time.perf_counter()
        run_stream(steps)
        torch.cuda.synchronize()
        stream_s = c.max_over_ranks(time.perf_counter() - t0)
        stream_fps = c.world * B * steps / stream_s
        rows_per_frame = kept_rows_per_frame(ex, cfg)
        h2d = B * W * H
        d2h = B * min(rows_per_frame, cap) * 60 + B * 4 + (pairs * (min(rows_per_frame, cap) * 8 + 4) if stereo else 0)
        api = ("orbx_stereo_batch_submit/_wait" if stereo else "orbx_extract_batch_submit/_wait")
        e2e_obj = {"value": stream_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "frames_per_step": B, "gpu_launches_per_step": per_step_launches,
                   "api": api + " (pinned host frames in, pinned host keypoints + descriptors" +
                          (" + mvuRight / mvDepth" if stereo else "") + " out; the steps of a continuous sequence, two in "
                          "flight, chunks rotating over 4 streams; every step's H2D and D2H inside the timed region)"}
        if with_sync_call:
            k = min(steps, 5)
            h_n = np.zeros(B, np.int32)
            for _ in range(2):
                ex.extract_batch_into(hnp, h_sets[0][0], h_sets[0][1], h_n)
            c.barrier()
            t0 = time.perf_counter()
            for _ in range(k):
                ex.extract_batch_into(hnp, h_sets[0][0], h_sets[0][1], h_n)   # synchronous: results are in host memory on return
            sync_s = c.max_over_ranks(time.perf_counter() - t0)
            e2e_obj["synchronous_call"] = {"value": c.world * B * k / sync_s, "unit": "frames/s", "steps": k,
                                           "api": "orbx_extract_batch, one blocking call per step: pays one chunk of "
                                                  "pipeline fill and drain per call"}
        # the box's copy ceiling with all ranks copying at once, in the same run
        pc = pcie_probe(c, h2d, max(d2h, 1 << 20))
        ceiling_fps = pc["h2d_gbs_with_d2h"] * 1e9 / (W * H)
        e2e_obj["pcie_ceiling_gbs"] = pc["h2d_gbs_with_d2h"]
        e2e_obj["pcie"] = pc
        e2e_obj["achieved_h2d_gbs"] = stream_fps * W * H / 1e9
        e2e_obj["frac_of_ceiling"] = stream_fps / ceiling_fps if ceiling_fps > 0 else None

    # ---- parity, outside the timed regions: sampled frames (and stereo pairs) of THIS rank's batch vs the oracle;
    # both legs must agree with each other on every frame
    parity = None
    if not args.skip_parity:
        kps_dev = unpack_kps(pkg, d_kps)
        desc_dev = d_desc.cpu().numpy()
        sample = sorted({0, 1, B // 2 & ~1, (B // 2 & ~1) + 1, B - 2, B - 1}) if stereo else sorted({0, B // 3, 2 * B // 3, B - 1})
        fr_host = {i: frames[i].cpu().numpy() for i in sample}
        got = {i: (kps_dev[i, :max(n_dev[i], 0)], desc_dev[i, :max(n_dev[i], 0)]) for i in sample}
        pr = None
        if stereo:
            ur, z, nm = d_ur.cpu().numpy(), d_z.cpu().numpy(), d_nm.cpu().numpy()
            pr = {(i, i + 1): (ur[i // 2], z[i // 2], int(nm[i // 2])) for i in sample if i % 2 == 0}
        parity = oracle_check(pkg, cfg, fr_host, got, pr)
        parity["legs_agree"] = None
        if h_sets is not None:
            dig_dev = frame_digests(n_dev, kps_dev, desc_dev, *((ur, z, nm) if stereo else ()))
            agree = True
            for s in h_sets:
                agree = agree and np.array_equal(frame_digests(s[2], s[0], s[1], *(s[3:6] if stereo else ())), dig_dev)
            parity["legs_agree"] = bool(agree)
        ok = not parity["mismatches"] and parity["legs_agree"] is not False and overflowed == 0
        ok = c.min_over_ranks(1.0 if ok else 0.0) == 1.0
        parity["result"] = "bit-exact" if ok else "MISMATCH"
        parity["what"] = (f"{len(sample)} frames" + (f" + {len(pr)} stereo pairs" if pr else "") + " per rank of the timed batch vs "
                          "the CPU oracle (keypoints incl. angle, descriptors" + (", mvuRight, mvDepth" if pr else "") +
                          "); every frame of the e2e leg vs the device-resident leg by digest; all ranks must agree")

    # ---- roofline of the dominant stage, from the event times of the timed region ----
    ab = algorithmic_bytes(cfg)
    stages = {}
    for k in ("pyramid", "fast", "octree", "blur", "describe"):
        t = stage_acc[k] / steps
        gbs = ab[k] * B / (t * 1e-3) / 1e9 if t > 0 else 0.0
        stages[k] = {"ms_per_step": t, "algorithmic_bytes_per_step": ab[k] * B, "gbs": gbs, "frac_of_hbm": gbs / c.peak}
    dom = max(("pyramid", "fast", "octree", "blur", "describe"), key=lambda k: stages[k]["ms_per_step"])
    kernel_names = {"pyramid": "resize4_kernel (7 launches)", "fast": "fast_cells_kernel", "octree": "octree_kernel",
                    "blur": "blur_kernel", "describe": "describe_kernel"}
    traffic, limiter = ncu_digest_for(kernel_names[dom], B, name)
    roof = {"bound": "hbm", "kernel": kernel_names[dom], "achieved": stages[dom]["gbs"], "peak": c.peak, "unit": "GB/s",
            "frac": stages[dom]["gbs"] / c.peak, "traffic": traffic, "limiter": limiter, "peak_source": c.peak_src,
            "share_of_step": stages[dom]["ms_per_step"] / (stage_acc["total"] / steps),
            "pipeline_achieved": ab["total"] * value / c.world / 1e9,
            "pipeline_frac": ab["total"] * value / c.world / 1e9 / c.peak,
            "algorithmic_bytes_per_frame": ab["total"]}
    out = {"metric": cfg["metric"], "value": value, "unit": "frames/s", "ms_per_step": ms / steps, "steps": steps,
           "config": static_config(cfg), "keypoints_per_frame": nkp / B, "overflowed_frames": overflowed,
           "gpu_launches": launches, "e2e": e2e_obj, "roofline": roof, "stages": stages, "parity": parity, "clocks": clocks}
    if stereo:
        out["stereo"] = {"pairs_per_s": value / 2, "matched_per_pair": float(d_nm.float().mean().item()),
                         "note": "value counts frames; every timed step also runs orbx_stereo_match_batch_device on its "
                                 f"{pairs} pairs (3 launches: row table, match + SAD refinement, median filter)"}
    out["_frames_host"] = [f for f in frames[:BASE_FRAMES].cpu().numpy()]    # for the CPU baseline leg; removed before printing
    return out


def csrc_hash():
    h = hashlib.sha256()
    # the sources of the extraction-pipeline kernels (the ones profiles/ncu_digest.json describes)
    for name in ("orb_pattern.inc", "orbx_common.cuh", "orbx_describe.cu", "orbx_fast.cu", "orbx_math.cuh", "orbx_octree.cu",
                 "orbx_pyramid.cu"):
        h.update(name.encode())
        h.update((ROOT / "orb-slam2-_b200" / "csrc" / name).read_bytes())
    return h.hexdigest()[:16]


def ncu_digest_for(kernel, frames_per_launch, workload):
    """DRAM traffic and the limiting pipe of a kernel from the committed ncu --set full capture (per launch, like
    `achieved`) -- only when the digest was made from exactly the kernel sources being timed (hash of csrc/*.cu*) for
    this workload and launch size; otherwise null."""
    p = ROOT / "profiles" / "ncu_digest.json"
    if not p.exists():
        return None, None
    dg = json.loads(p.read_text())
    k = dg.get("kernels", {}).get(kernel)
    if not k or dg.get("frames_per_launch") != frames_per_launch or dg.get("workload", "C1") != workload:
        return None, None
    if dg.get("csrc_sha16") != csrc_hash():
        return None, {"stale": f"profiles/ncu_digest.json describes csrc {dg.get('csrc_sha16')}, not the kernels timed here"}
    return k["dram_bytes"], {"pipe": "alu (integer issue)", "pct_of_peak": k.get("alu_pipe_pct"), "issue_pct": k.get("issue_pct"),
                             "source": dg["source"]}


def run_sequence_c3(c, pkg, args):
    """BASELINE.json config 3 as specified: ONE EuRoC-shaped stereo sequence of 10 000 distinct 752x480 frames (5 000
    L/R pairs), sharded pair-aligned over the ranks (orbx_shard_extract_sequence: rank r streams its block through
    orbx_stereo_batch_submit/_wait from pinned HOST memory, both extractions + Frame::ComputeStereoMatches per pair,
    results back in pinned host arrays). The frames are generated on the device once, copied to the host before the
    timed region, and are a function of the frame index only, so the digest of the whole sequence's results is the
    same at every N."""
    torch = c.torch
    cfg = WORKLOADS["C3"]
    W, H, NFEAT, total = cfg["w"], cfg["h"], cfg["nfeat"], cfg["sequence"]
    lo, hi = pkg.sharding.frame_range(total, c.rank, c.world, group=2)
    cnt = hi - lo
    base = base_frames(pkg, cfg)
    hnp = pkg.pinned_empty((cnt, H, W), np.uint8, device=c.local)
    host = torch.from_numpy(hnp)
    piece = 512
    for f0 in range(0, cnt, piece):
        m = min(piece, cnt - f0)
        host[f0:f0 + m].copy_(device_frames(torch, base, lo + f0, m, salt=0))
    torch.cuda.synchronize()
    cap = NFEAT + 64
    kps = pkg.pinned_empty((cnt, cap), pkg.KP_DTYPE, device=c.local)
    desc = pkg.pinned_empty((cnt, cap, 32), np.uint8, device=c.local)
    n = pkg.pinned_empty(cnt, np.int32, device=c.local)
    ur = pkg.pinned_empty((cnt // 2, cap), np.float32, device=c.local)
    dp = pkg.pinned_empty((cnt // 2, cap), np.float32, device=c.local)
    nm = pkg.pinned_empty(cnt // 2, np.int32, device=c.local)
    params = dict(nfeatures=NFEAT, scale_factor=SCALE, nlevels=NLEVELS, ini_th_fast=INI_TH, min_th_fast=MIN_TH,
                  max_width=W, max_height=H, max_batch=256)
    g = c.shard_group
    st = (cfg["mbf"], 0.0, ur, dp, nm)

    def one_pass():
        g.extract_sequence(params, hnp, kps, desc, n, group=2, stereo=st, total=total, block=(lo, hi))

    one_pass()                                   # warm-up pass (lanes, arenas, pinned staging)
    passes = max(1, min(args.steps, 3))
    l0 = g.launch_count()
    c.barrier()
    t0 = time.perf_counter()
    for _ in range(passes):
        one_pass()
    torch.cuda.synchronize()
    dt = c.max_over_ranks(time.perf_counter() - t0)
    launches = g.launch_count() - l0
    fps = total * passes / dt
    # parity: every frame's digest, gathered in sequence order -> one digest of the whole sequence (the same at every
    # N), plus sampled frames / pairs of this rank's block against the oracle
    dig = frame_digests(n, kps, desc, ur, dp, nm)
    if c.world > 1:
        allg = [None] * c.world
        c.dist.all_gather_object(allg, dig.tobytes())
        seq_bytes = b"".join(allg)
    else:
        seq_bytes = dig.tobytes()
    seq_digest = hashlib.sha256(seq_bytes).hexdigest()[:32]
    parity = None
    if not args.skip_parity:
        sample = sorted({0, 1, (cnt // 2) & ~1, ((cnt // 2) & ~1) + 1, cnt - 2, cnt - 1})
        fr_host = {i: hnp[i] for i in sample}
        got = {i: (kps[i, :n[i]], desc[i, :n[i]]) for i in sample}
        pr = {(i, i + 1): (ur[i // 2], dp[i // 2], int(nm[i // 2])) for i in sample if i % 2 == 0}
        parity = oracle_check(pkg, cfg, fr_host, got, pr)
        ok = not parity["mismatches"] and int((n < 0).sum()) == 0 and len(seq_bytes) == total * 8
        ok = c.min_over_ranks(1.0 if ok else 0.0) == 1.0
        parity["result"] = "bit-exact" if ok else "MISMATCH"
        parity["sequence_digest"] = seq_digest
        parity["what"] = ("3 stereo pairs (6 frames) of every rank's block vs the CPU oracle (keypoints, descriptors, mvuRight, "
                          "mvDepth); sequence_digest = sha256 over the per-frame digests of all 10000 frames in sequence "
                          "order -- a function of the frames only, hence identical at N = 1, 2, 4, 8")
    rows = kept_rows_per_frame(pkg.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=c.local, max_width=W, max_height=H), cfg)
    return {"value": fps, "unit": "frames/s", "frames": total, "pairs": total // 2, "passes": passes,
            "ms_per_pass": dt / passes * 1e3, "pairs_per_s": fps / 2, "frames_this_rank": cnt,
            "h2d_bytes_per_pass": total * W * H, "d2h_bytes_per_pass": total * (min(rows, cap) * 60 + 4) + (total // 2) * (min(rows, cap) * 8 + 4),
            "achieved_h2d_gbs": fps * W * H / 1e9, "gpu_launches_per_pass": int(c.sum_over_ranks(launches)) // passes,
            "matched_per_pair": float(nm.mean()), "keypoints_per_frame": float(n.mean()),
            "frames_source": "generated on the device before the timed region (a function of the frame index), copied to "
                             "pinned host memory; the timed passes stream them from the HOST",
            "api": "orbx_shard_extract_sequence -> orbx_stereo_batch_submit/_wait per rank (256-frame submissions, two in "
                   "flight, chunks over 4 streams), H2D and D2H inside the timed region",
            "parity": parity}


def ham_block(torch, dev, b):
    """Rows [b*HAM_BLOCK, (b+1)*HAM_BLOCK) of the config-5 database: uniform random bytes, seeded by the block index."""
    g = torch.Generator(device=dev)
    g.manual_seed(5000 + b)
    return torch.randint(0, 256, (HAM_BLOCK, 32), generator=g, device=dev, dtype=torch.uint8)


def run_hamming(c, pkg, args):
    """Second metric of BASELINE.json ("Hamming match Gpairs/s"), config 5: 10k queries against a 10M-row descriptor
    database split into `world` contiguous row ranges (strong scaling: the database is fixed), queries replicated,
    through orbx_shard_match_bf2_device: per-rank top-2, ncclAllGather (issued by liborbx.so) of 12 B per query and
    rank, merge + ratio test on every rank. Timed on the device, max over ranks; database and queries resident in HBM.
    Inputs per SURVEY 8d: uniform random bytes; 5000 queries are database rows with k in U[0,40] flipped bits."""
    torch = c.torch
    dev = torch.device("cuda", c.local)
    world, rank = c.world, c.rank
    lo, hi = pkg.sharding.block_range(HAM_NDB, rank, world)
    assert lo % HAM_BLOCK == 0 and hi % HAM_BLOCK == 0
    db = torch.cat([ham_block(torch, dev, b) for b in range(lo // HAM_BLOCK, hi // HAM_BLOCK)])
    # queries: seed 77, the same on every rank; query i < 5000 is planted from global row plant_row[i]
    rng = np.random.Generator(np.random.PCG64(77))
    q = rng.integers(0, 256, (HAM_NQ, 32), dtype=np.uint8)
    # 5000 plants with k in U[0,40] flipped bits (all pass TH_LOW = 50 and, the second-best of 10M random
    # rows being ~85, the 0.6 ratio) + 1000 with k in U[41,72], which fail TH_LOW or the ratio: both branches of
    # R/src/ORBmatcher.cpp:554-556 are taken at scale
    n40, nfar = HAM_NQ // 2, HAM_NQ // 10
    nplant = n40 + nfar
    plant_row = np.sort(rng.choice(HAM_NDB, nplant, replace=False))
    plant_k = np.concatenate([rng.integers(0, 41, n40), rng.integers(41, 73, nfar)])
    flips = np.zeros((nplant, 32), np.uint8)
    for i in range(nplant):
        bits = rng.choice(256, int(plant_k[i]), replace=False)
        np.bitwise_or.at(flips[i], bits >> 3, (1 << (bits & 7)).astype(np.uint8))
    # the planted rows' bytes: regenerate only the blocks that hold them is 80 blocks at most -> every rank asks the
    # owners instead: each rank contributes the rows it owns, summed over the ranks (others contribute zeros)
    mine = (plant_row >= lo) & (plant_row < hi)
    rows = torch.zeros((nplant, 32), dtype=torch.int32, device=dev)
    rows[torch.from_numpy(np.nonzero(mine)[0]).to(dev)] = db[torch.from_numpy(plant_row[mine] - lo).to(dev)].to(torch.int32)
    if world > 1:
        c.dist.all_reduce(rows)
    q[:nplant] = rows.cpu().numpy().astype(np.uint8) ^ flips
    dq = torch.from_numpy(q).to(dev)
    g = c.shard_group
    g.set_database_device(0, db.data_ptr(), lo, hi, HAM_NDB)
    out = torch.empty((3, HAM_NQ), dtype=torch.int32, device=dev)
    acc = torch.empty(HAM_NQ, dtype=torch.uint8, device=dev)

    def step():
        g.match_bf2_device(0, dq.data_ptr(), HAM_NQ, 0.6, 50, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                           acc.data_ptr(), c.stream)

    for _ in range(3):
        step()
    steps = max(1, min(args.steps, 5))
    sampler = ClockSampler(c.local)
    l0 = g.launch_count()
    c.barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    c.barrier()
    ms = c.max_over_ranks(e0.elapsed_time(e1)) / steps
    clocks = sampler.stop() if rank == 0 else None
    launches = (g.launch_count() - l0) // steps
    pairs = float(HAM_NQ) * float(HAM_NDB)
    gp = pairs / (ms * 1e-3) / 1e9

    # ---- parity (outside the timed region) ----
    parity = None
    if not args.skip_parity:
        o = out.cpu().numpy()
        a = acc.cpu().numpy()
        # (1) every planted query returns its planted global row at distance k; accepted iff k <= 50 and k < 0.6 * d2
        p_ok = np.array_equal(o[0][:nplant], plant_row.astype(np.int32)) and np.array_equal(o[1][:nplant], plant_k.astype(np.int32))
        exp_acc = (o[1] <= 50) & (o[1].astype(np.float32) < np.float32(0.6) * o[2].astype(np.float32))
        p_ok = p_ok and np.array_equal(a.astype(bool), exp_acc) and int(a[nplant:].sum()) == 0
        p_ok = p_ok and int(a[:n40].sum()) == n40 and 0 < int(a[n40:nplant].sum()) < nfar
        # (2) the N-rank answer equals the 1-rank answer on the first 1M rows: every rank rebuilds the slice, runs the
        # single-GPU entry points on all of it and the sharded path on its part of it
        S = 1_000_000
        sl = torch.cat([ham_block(torch, dev, b) for b in range(S // HAM_BLOCK)])
        m = pkg.ORBmatcher(0.6, device=c.local)
        part = torch.empty((3, HAM_NQ), dtype=torch.int32, device=dev)
        one = torch.empty((3, HAM_NQ), dtype=torch.int32, device=dev)
        one_acc = torch.empty(HAM_NQ, dtype=torch.uint8, device=dev)
        m.match_bf2_device(dq.data_ptr(), HAM_NQ, sl.data_ptr(), S, 0, part[0].data_ptr(), part[1].data_ptr(), part[2].data_ptr(), c.stream)
        m.merge_device(part[0].data_ptr(), part[1].data_ptr(), part[2].data_ptr(), 1, HAM_NQ, 50, one[0].data_ptr(),
                       one[1].data_ptr(), one[2].data_ptr(), one_acc.data_ptr(), c.stream)
        slo, shi = pkg.sharding.block_range(S, rank, world)
        g.set_database_device(0, sl[slo:shi].data_ptr(), slo, shi, S)
        sh = torch.empty((3, HAM_NQ), dtype=torch.int32, device=dev)
        sh_acc = torch.empty(HAM_NQ, dtype=torch.uint8, device=dev)
        g.match_bf2_device(0, dq.data_ptr(), HAM_NQ, 0.6, 50, sh[0].data_ptr(), sh[1].data_ptr(), sh[2].data_ptr(),
                           sh_acc.data_ptr(), c.stream)
        torch.cuda.synchronize()
        n_ok = bool(torch.equal(sh, one) and torch.equal(sh_acc, one_acc))
        g.set_database_device(0, db.data_ptr(), lo, hi, HAM_NDB)
        # (3) 64 queries x the 1M-row slice against the CPU oracle (rank 0)
        o_ok = True
        if rank == 0:
            import oracle_lib as ol
            sel = np.r_[0:24, n40:n40 + 16, nplant:nplant + 24]
            e = ol.match_bf2(q[sel], sl.cpu().numpy(), np.float32(0.6), 50, nthreads=os.cpu_count() or 1)
            oo, oa = one.cpu().numpy(), one_acc.cpu().numpy()
            o_ok = all(np.array_equal(oo[k][sel], e[k]) for k in range(3)) and np.array_equal(oa[sel], e[3])
        ok = c.min_over_ranks(1.0 if (p_ok and n_ok and o_ok) else 0.0) == 1.0
        parity = {"result": "bit-exact" if ok else "MISMATCH", "planted_rows_and_distances": bool(p_ok),
                  "n_rank_equals_1_rank_on_1M_rows": n_ok, "oracle_64_queries_x_1M_rows": bool(o_ok),
                  "accepted": int(a.sum()), "planted": nplant, "plant_seed": 77,
                  "what": "5000 planted queries with k in U[0,40] flipped bits (all accepted) and 1000 with k in U[41,72] (rejected by "
                          "TH_LOW or by the ratio, a few accepted) return their planted global row with d1 = k, and `accepted` equals "
                          "the ratio / TH_LOW decision recomputed on the host; (idx, d1, d2, accepted) of the sharded path == the "
                          "single-GPU entry points on a 1M-row slice; 64 queries of it == the CPU oracle"}
        del sl

    # ---- the bound: integer issue, not HBM ----
    clk = (clocks or {}).get("sm_mhz") or 1965.0
    sms = torch.cuda.get_device_properties(c.local).multi_processor_count
    ALU_OPS = 16.75   # 8 XOR + 8 LOP3 (carry-save adders) + 0.5 VIMNMX3 + 0.25 ISETP per pair (csrc/orbx_match.cu)
    alu_peak = sms * 64 * clk * 1e6 / ALU_OPS / 1e9 * world    # Gpairs/s if the ALU pipe did nothing but this kernel's ops
    popc_peak = sms * 16 * clk * 1e6 / 4 / 1e9 * world
    roof = {"bound": "alu", "kernel": "bf2_kernel", "achieved": gp, "peak": alu_peak, "unit": "Gpairs/s", "frac": gp / alu_peak,
            "traffic": None,
            "peak_source": f"{sms} SMs x 64 ALU-pipe lanes/clk x {clk:.0f} MHz (median under load) / {ALU_OPS} ALU-pipe ops per pair "
                           "(8 XOR + 8 LOP3 + 0.5 VIMNMX3 + 0.25 ISETP; pipe rates: profiles/r1_pipe_probe.txt), x N GPUs; the 4 POPC per pair "
                           "on the quarter-rate pipe (popc_pipe_peak) co-limit the loop",
            "popc_pipe_peak": popc_peak, "popc_pipe_frac": gp / popc_peak,
            "hbm_gbs": (HAM_NDB * 32 / world) / (ms * 1e-3) / 1e9, "hbm_frac": (HAM_NDB * 32 / world) / (ms * 1e-3) / 1e9 / c.peak,
            "note": "not HBM-bound by construction: 320 MB of compulsory traffic against 1e11 pairs"}
    res = {"metric": "Hamming match Gpairs/s", "value": gp, "unit": "Gpairs/s", "ms_per_step": ms, "steps": steps,
           "config": {"workload": f"C5: brute-force top-2 + ratio, {HAM_NQ} queries x {HAM_NDB} database rows (32 B), "
                                  f"database sharded x{world}, ncclAllGather + merge inside liborbx.so (orbx_shard_match_bf2_device)",
                      "scaling": "strong"},
           "gpu_launches_per_step": launches, "nccl": g.info(),
           "compulsory_bytes_per_step": HAM_NDB * 32 + HAM_NQ * 32 * world + HAM_NQ * 13,
           "roofline": roof, "parity": parity, "clocks": clocks, "cpu_baseline": None}
    if world == 1 and rank == 0 and not args.skip_cpu:
        res["cpu_baseline"] = hamming_cpu_baseline(pkg, q, torch.cat([ham_block(torch, dev, b) for b in range(8)]).cpu().numpy())
    return res


def run_matcher_rows(pkg, local):
    """SURVEY 8f rows (key-frame searches, distinctive descriptors, frame epilogue, vocabulary descent): ms per call
    through the product's host-buffer entry points next to the oracle port on ONE host thread, same inputs. These
    are per-frame latencies of small problems (10^3 keypoints), reported for completeness; rank 0, N = 1 only."""
    import oracle_lib as ol
    import tempfile
    import voc_case

    def ms(fn, reps):
        fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        return (time.perf_counter() - t0) / reps * 1e3

    rng = np.random.default_rng(5)
    w, h = 640, 480
    ex = pkg.ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h, device=local)
    m = pkg.ORBmatcher(0.6, device=local)
    img = pkg.synth.frame(w, h, 70)
    k1, d1 = ex(img)
    k2, d2 = ex(np.roll(img, (3, 6), (0, 1)))
    rows = {}
    node = lambda d: ((d[:, 0] >> 4).astype(np.int32) << 3) | (d[:, 5] >> 5).astype(np.int32)
    blocks = ol.common_blocks(ol.featvec(node(d1)), ol.featvec(node(d2)))
    pairs = int(np.sum(np.diff(blocks[0]).astype(np.int64) * np.diff(blocks[2])))
    rows["bow_blocks"] = {"what": f"orbx_match_blocks, {len(k1)} x {len(k2)} features in {len(blocks[0]) - 1} shared nodes, {pairs} pairs",
                          "gpu_ms": ms(lambda: m.match_blocks(d1, d2, *blocks), 50),
                          "cpu_ms": ms(lambda: ol.block_distances(d1, d2, *blocks), 20)}
    sizes = rng.integers(2, 13, 2000).astype(np.int32)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    obs = pkg.synth.descriptors(int(off[-1]), 4)
    rows["distinctive"] = {"what": f"orbx_distinctive, 2000 map points, {int(off[-1])} observations",
                           "gpu_ms": ms(lambda: m.distinctive(obs, off), 50), "cpu_ms": ms(lambda: ol.distinctive(obs, off), 10)}
    K4 = np.array([517.3, 516.5, 318.6, 255.3], np.float32)
    D = np.array([0.2624, -0.9531, -0.0054, 0.0026, 1.1633], np.float32)
    xy = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    rows["undistort"] = {"what": f"orbx_undistort_points, {len(xy)} keypoints, 5 coefficients",
                         "gpu_ms": ms(lambda: pkg.undistort_points(xy, K4, D, device=local), 50),
                         "cpu_ms": ms(lambda: ol.undistort_points(xy, K4, D), 50)}
    fr = pkg.Frame(m)
    b = np.array([0, w, 0, h], np.float32)
    fr.set(k2, d2, b)
    nq = 1000
    i = rng.integers(0, len(k1), nq)
    q = (xy[i], np.full(nq, 15, np.float32), np.maximum(k1["octave"][i] - 1, 0).astype(np.int32),
         (k1["octave"][i] + 1).astype(np.int32), d1[i].copy())
    grid = ol.Grid(k2, b)

    def cpu_window():
        for j in range(nq):
            cnd = grid.query(q[0][j, 0], q[0][j, 1], q[1][j], int(q[2][j]), int(q[3][j]))
            np.unpackbits(d2[cnd] ^ q[4][j], axis=1).sum(1)
    rows["window_search"] = {"what": f"orbx_frame_match_window, {nq} queries on a resident {len(k2)}-keypoint frame, radius 15",
                             "gpu_ms": ms(lambda: fr.match_window(*q), 50), "cpu_ms": ms(cpu_window, 2),
                             "cpu_note": "oracle grid query per point + numpy popcount (Python loop)"}
    with tempfile.TemporaryDirectory() as td:
        path = Path(td) / "voc.txt"
        voc_case.write(path, 10, 5, 0, 0, seed=3, min_leaf_level=4)
        ov = ol.Vocabulary(path)
        parent, ndesc, weight, word, klsw = ov.export()
        gv = pkg.Vocabulary(m, parent, ndesc, int(klsw[1]))
        rows["vocabulary"] = {"what": f"orbx_vocabulary_descend, {len(d1)} descriptors, {len(parent)} nodes (k 10, L 5)",
                              "gpu_ms": ms(lambda: gv.descend(d1, 3), 50), "cpu_ms": ms(lambda: ov.transform(d1, 3), 10),
                              "cpu_note": "oracle transform incl. BowVector / FeatureVector assembly"}
    return rows


def run_gpu(args):
    import torch
    # The C ABI treats a NULL stream as "the handle's own stream"; torch's default stream has handle 0, so all
    # device work of this process runs on one explicit side stream: the kernels, NCCL and the timing events
    # are then ordered on the same stream.
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; liborbx has no CPU fallback")
    torch.cuda.set_device(local)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        return _run_gpu(args, torch)


def _run_gpu(args, torch):
    import orbslam2_b200 as pkg
    c = make_ctx(torch)
    # the product's multi-GPU component (csrc/orbx_shard.cu): one rank per process, NCCL bound by liborbx.so; the id
    # travels through the launcher's own process group
    uid = None
    if c.world > 1:
        t = torch.from_numpy(pkg.sharding.unique_id()).cuda() if c.rank == 0 else torch.zeros(128, dtype=torch.uint8, device="cuda")
        c.dist.broadcast(t, 0)
        uid = t.cpu().numpy()
    c.shard_group = pkg.sharding.ShardGroup.from_rank(c.local, c.rank, c.world, uid)

    top = run_extraction(c, pkg, args.workload, args, args.steps, with_sync_call=True)
    frames_host = top.pop("_frames_host")
    sub_steps = max(1, min(args.steps, 5))
    workloads = {}
    if not args.skip_workloads:
        for name in ("C2", "C3", "C4"):
            if name == args.workload:
                continue
            w = run_extraction(c, pkg, name, args, sub_steps)
            w.pop("_frames_host")
            if name == "C3":
                w["sequence"] = run_sequence_c3(c, pkg, args)
            workloads[name] = w
            torch.cuda.empty_cache()
    hamming = None if args.skip_hamming else run_hamming(c, pkg, args)

    if c.rank != 0:
        c.shard_group.close()
        if c.world > 1:
            c.dist.destroy_process_group()
        return 0

    cfg = WORKLOADS[args.workload]
    cpu = None
    if c.world == 1 and not args.skip_cpu:
        cores = os.cpu_count() or 1
        fps, n, med_ms = time_cpu_extractor(cfg, frames_host, 12.0, cores)
        cpu = {"value": fps, "unit": "frames/s", "cores": cores, "kind": cpu_extractor_factory(cfg)[1],
               "sample": f"12 s of the same {cfg['w']}x{cfg['h']} frames on {cores} threads ({n} frames), one extractor per thread",
               "single_thread_ms_per_frame": med_ms, "note": CPU_NOTE,
               "opencv_primitives_ms": opencv_primitives_ms(cfg, frames_host)}
    matcher_rows = run_matcher_rows(pkg, c.local) if (c.world == 1 and not args.skip_cpu and not args.skip_hamming) else None

    line = {"metric": top["metric"], "value": top["value"], "unit": "frames/s", "n_gpus": c.world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": top["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": top["config"], "keypoints_per_frame": top["keypoints_per_frame"],
            "overflowed_frames": top["overflowed_frames"], "clocks": top["clocks"], "e2e": top["e2e"],
            "gpu_launches": top["gpu_launches"], "roofline": top["roofline"], "stages": top["stages"],
            "parity": top["parity"], "stereo": top.get("stereo"), "cpu_baseline": cpu, "workloads": workloads,
            "hamming": hamming, "matcher_rows": matcher_rows, "csrc_sha16": csrc_hash()}
    print(json.dumps(line), flush=True)
    c.shard_group.close()
    if c.world > 1:
        c.dist.destroy_process_group()
    return 0


def main():
    # NCCL / torchrun may write to stdout; the contract is ONE JSON line there. Everything else goes to stderr.
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="orbx", choices=["orbx", "reference"])
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs only: skip the host-buffer leg")
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs only: skip the CPU baseline legs")
    ap.add_argument("--skip-hamming", action="store_true", help="profiling runs only: skip the Hamming (config 5) leg")
    ap.add_argument("--skip-workloads", action="store_true", help="profiling runs only: skip configs 2-4")
    ap.add_argument("--skip-parity", action="store_true", help="profiling runs only: skip the oracle comparisons")
    ap.add_argument("--workload", default="C1", choices=sorted(WORKLOADS), help="top-level extraction config (default: the metric's C1)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
