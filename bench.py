#!/usr/bin/env python3
"""bench.py -- the driver's measurement contract for the ORB front-end hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json `metric`): ORBextractor::operator() on 640x480 u8 frames, 1000 features,
8 levels, scale 1.2, FAST 20/7 (TUM1.yaml; config C1 of SURVEY.md section 8). One "step" = one pass
of the extraction pipeline over a batch of FRAMES_PER_STEP distinct synthetic frames (larger than
the 126 MB L2). Frames shard across ranks with no data-path collective ("weak" scaling: the
per-GPU batch is fixed).

  value : frames/s, inputs and outputs resident in HBM (orbx_extract_batch_device)
  e2e   : frames/s through the host-buffer C-ABI call a drop-in ORBextractor makes
          (orbx_extract_batch: pinned host frames -> H2D -> kernels -> D2H keypoints+descriptors)
  roofline / stages : per-stage CUDA-event times from inside the timed region
  cpu_baseline : the CPU oracle on this box's host cores, bounded sample (rank 0, N=1 only)

--impl reference times the CPU restatement of the reference (oracle/) with all host threads on the
same config; this is the only place outside tests/ and smoke() where oracle/ is executed.
"""
import argparse
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

W, H, NFEAT, NLEVELS, SCALE, INI_TH, MIN_TH = 640, 480, 1000, 8, 1.2, 20, 7
FRAMES_PER_STEP = 512          # 512 x 307200 B = 157 MB of input per step (> 126 MB L2)
E2E_FRAMES_PER_STEP = 512
BASE_FRAMES = 32               # distinct numpy-generated frames; the rest are shifted/noised copies
METRIC = "ORB frames/s (640x480, 1000 kp)"
WORKLOAD = ("C1: ORBextractor::operator() on 640x480 u8 frames, 1000 features, 8 levels, scale 1.2, "
            "FAST 20/7 (TUM1.yaml)")


# The other extraction shapes BASELINE.json names (parity-test cases; `--workload` reports their throughput too,
# the default and the driver's line stay C1). Frames per step keep ~150 MB of distinct input per step.
WORKLOADS = {
    "C1": (640, 480, 1000, 512, METRIC, WORKLOAD),
    "C2": (1241, 376, 2000, 320, "ORB frames/s (1241x376, 2000 kp)",
           "C2 shape: KITTI 1241x376 stereo images, 2000 features per image (left and right frames are independent frames "
           "of the batch; row matching is orbx_stereo_match, timed in tools/match_probe.py)"),
    "C3": (752, 480, 1200, 416, "ORB frames/s (752x480, 1200 kp)", "C3 shape: EuRoC 752x480 frames, 1200 features, sharded by frames"),
    "C4": (1920, 1080, 4000, 72, "ORB frames/s (1920x1080, 4000 kp)", "C4: 1920x1080 frames, 4000 features, 8 levels, batched"),
}


def select_workload(name):
    global W, H, NFEAT, FRAMES_PER_STEP, E2E_FRAMES_PER_STEP, METRIC, WORKLOAD
    W, H, NFEAT, FRAMES_PER_STEP, METRIC, WORKLOAD = WORKLOADS[name]
    E2E_FRAMES_PER_STEP = FRAMES_PER_STEP


def level_sizes():
    inv = [np.float32(1.0)]
    sc = [np.float32(1.0)]
    for _ in range(1, NLEVELS):
        sc.append(np.float32(np.float64(sc[-1]) * np.float64(np.float32(SCALE))))
    inv = [np.float32(1.0) / s for s in sc]
    return [(int(np.rint(np.float32(W) * s)), int(np.rint(np.float32(H) * s))) for s in inv]


def algorithmic_bytes():
    """bytes per frame, per stage and total (5 702 474 B at C1)."""
    A = [w * h for w, h in level_sizes()]
    sA = sum(A)
    st = {
        "pyramid": (sA - A[-1]) + (sA - A[0]),
        "fast": sA,
        "blur": 2 * sA,
        "describe": NFEAT * (709 + 512) + NFEAT * 60,
        "octree": 0,
    }
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


def base_frames(pkg, n=BASE_FRAMES, seed0=1000):
    if METRIC.startswith("ORB frames/s (1241x376"):   # C2: stereo pairs, interleaved L R L R ...
        return np.stack([im for i in range(n // 2) for im in pkg.synth.stereo_pair(W, H, seed0 + i)])
    return np.stack([pkg.synth.frame(W, H, seed0 + i) for i in range(n)])


def device_frames(torch, base, count, rank):
    """count distinct frames on the device: base frames rolled by a per-copy offset plus +-1 noise
    (deterministic torch generator). Generation happens outside every timed region."""
    dev = torch.device("cuda", torch.cuda.current_device())
    b = torch.from_numpy(base).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    out = torch.empty((count, H, W), dtype=torch.uint8, device=dev)
    nb = b.shape[0]
    for k in range(0, count, nb):
        m = min(nb, count - k)
        rep = k // nb
        fr = torch.roll(b[:m], shifts=(3 * rep, 5 * rep), dims=(1, 2)).to(torch.int16)
        if rep:
            fr = fr + torch.randint(-1, 2, fr.shape, generator=g, device=dev, dtype=torch.int16)
        out[k:k + m] = fr.clamp_(0, 255).to(torch.uint8)
    return out


def cpu_extractor_factory():
    """(constructor, kind): the reference's own ORBextractor.cpp (oracle/_ref, built in the authoring container from
    /root/reference against the cv shim; the .so travels) when present -> kind "reference"; else the oracle port."""
    import ref_lib
    if ref_lib.available():
        ref_lib.lib()
        return (lambda: ref_lib.RefExtractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)), "reference"
    import oracle_lib
    oracle_lib.lib()
    return (lambda: oracle_lib.Extractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)), "port"


def time_cpu_oracle(frames, budget_s, nthreads):
    """frames/s of the CPU oracle with `nthreads` threads, each with its own extractor instance
    (the reference's own pattern: one ORBextractor per std::thread, R/src/Frame.cpp:86-89)."""
    nfr = len(frames)
    exs = [cpu_extractor_factory()[0]() for _ in range(nthreads)]
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


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; OpenCV C++ is not installable here, so
    the primitives are the cv2-4.13-pinned restatements) on all host threads, same config/metric."""
    import orbslam2_b200 as pkg  # synth only; no CUDA call is made on this path
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    frames = [pkg.synth.frame(W, H, 1000 + i) for i in range(BASE_FRAMES)]
    budget = 4.0
    for _ in range(min(args.warmup, 1)):
        time_cpu_oracle(frames, 1.0, cores)
    t0 = time.perf_counter()
    tot_frames, vals = 0, []
    for _ in range(args.steps):
        fps, n, med_ms = time_cpu_oracle(frames, budget, cores)
        vals.append(fps)
        tot_frames += n
    dt = time.perf_counter() - t0
    value = float(np.mean(vals))
    sample = f"{args.steps} steps x {budget:.0f} s of {W}x{H} frames on {cores} threads ({tot_frames} frames)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "frames_per_step": "time-bounded sample"},
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": cpu_extractor_factory()[1],
                             "sample": sample,
                             "single_thread_ms_per_frame": med_ms},
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def run_matcher_rows(pkg, local):
    """SURVEY 8f rows (key-frame searches, distinctive descriptors, frame epilogue, vocabulary descent): ms per call
    through the product's host-buffer entry points next to the oracle port on ONE host thread, same inputs. These
    are per-frame latencies of small problems (10^3 keypoints), reported for completeness; rank 0, N = 1 only."""
    import oracle_lib as ol
    import tempfile
    sys.path.insert(0, str(ROOT / "tests"))
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
    # node-restricted distance blocks of SearchByBoW / SearchForTriangulation (128 synthetic nodes)
    node = lambda d: ((d[:, 0] >> 4).astype(np.int32) << 3) | (d[:, 5] >> 5).astype(np.int32)
    blocks = ol.common_blocks(ol.featvec(node(d1)), ol.featvec(node(d2)))
    pairs = int(np.sum(np.diff(blocks[0]).astype(np.int64) * np.diff(blocks[2])))
    rows["bow_blocks"] = {"what": f"orbx_match_blocks, {len(k1)} x {len(k2)} features in {len(blocks[0]) - 1} shared nodes, {pairs} pairs",
                          "gpu_ms": ms(lambda: m.match_blocks(d1, d2, *blocks), 50),
                          "cpu_ms": ms(lambda: ol.block_distances(d1, d2, *blocks), 20)}
    # distinctive descriptors of 2000 map points with 2-12 observations
    sizes = rng.integers(2, 13, 2000).astype(np.int32)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    obs = pkg.synth.descriptors(int(off[-1]), 4)
    rows["distinctive"] = {"what": f"orbx_distinctive, 2000 map points, {int(off[-1])} observations",
                           "gpu_ms": ms(lambda: m.distinctive(obs, off), 50), "cpu_ms": ms(lambda: ol.distinctive(obs, off), 10)}
    # frame epilogue + windowed search on the resident frame (SearchByProjection's size)
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
            c = grid.query(q[0][j, 0], q[0][j, 1], q[1][j], int(q[2][j]), int(q[3][j]))
            np.unpackbits(d2[c] ^ q[4][j], axis=1).sum(1)
    rows["window_search"] = {"what": f"orbx_frame_match_window, {nq} queries on a resident {len(k2)}-keypoint frame, radius 15",
                             "gpu_ms": ms(lambda: fr.match_window(*q), 50), "cpu_ms": ms(cpu_window, 2),
                             "cpu_note": "oracle grid query per point + numpy popcount (Python loop)"}
    # vocabulary descent: k = 10, L = 5 (about 10^4 words), levelsup 3
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


HAM_NQ, HAM_NDB = 10_000, 10_000_000   # BASELINE.json config 5


def run_hamming(torch, dist, pkg, world, rank, local, args, barrier, max_over_ranks):
    """Second metric of BASELINE.json ("Hamming match Gpairs/s"), config 5: 10k queries against a 10M-row
    descriptor database split into `world` contiguous row ranges (strong scaling: the database is fixed),
    queries replicated, per-rank top-2, all-gather of (idx, d1, d2) over NCCL, merge + ratio test on every
    rank. Timed on the device, max over ranks; database and queries are resident in HBM."""
    dev = torch.device("cuda", local)
    g = torch.Generator(device=dev)
    g.manual_seed(99)
    q = torch.randint(0, 256, (HAM_NQ, 32), generator=g, device=dev, dtype=torch.uint8)   # same on every rank
    lo, hi = pkg.sharding.block_range(HAM_NDB, rank, world)
    g.manual_seed(1000 + rank)
    db = torch.randint(0, 256, (hi - lo, 32), generator=g, device=dev, dtype=torch.uint8)
    plant = torch.arange(rank, HAM_NQ // 2, world, device=dev)      # query i has ONE exact row, in shard i % world
    db[(plant // world) * ((hi - lo) // (HAM_NQ // 2 // world + 1))] = q[plant]
    m = pkg.ORBmatcher(0.6, device=local)
    part = torch.empty((3, HAM_NQ), dtype=torch.int32, device=dev)
    gathered = torch.empty((world, 3, HAM_NQ), dtype=torch.int32, device=dev)
    out = torch.empty((3, HAM_NQ), dtype=torch.int32, device=dev)
    acc = torch.empty(HAM_NQ, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    assert stream != 0

    def step():
        m.match_bf2_device(q.data_ptr(), HAM_NQ, db.data_ptr(), hi - lo, lo, part[0].data_ptr(), part[1].data_ptr(),
                           part[2].data_ptr(), stream)
        if world > 1:
            dist.all_gather_into_tensor(gathered, part)
            idx, d1, d2 = (gathered[:, k, :].contiguous() for k in range(3))
        else:
            idx, d1, d2 = part[0], part[1], part[2]
        m.merge_device(idx.data_ptr(), d1.data_ptr(), d2.data_ptr(), world, HAM_NQ, 50, out[0].data_ptr(),
                       out[1].data_ptr(), out[2].data_ptr(), acc.data_ptr(), stream)

    for _ in range(3):
        step()
    steps = max(1, min(args.steps, 5))
    l0 = m.launch_count()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    pairs = float(HAM_NQ) * float(HAM_NDB)
    gp = pairs / (ms * 1e-3) / 1e9
    return {"metric": "Hamming match Gpairs/s", "value": gp, "unit": "Gpairs/s", "ms_per_step": ms, "steps": steps,
            "config": {"workload": f"C5: brute-force top-2 + ratio, {HAM_NQ} queries x {HAM_NDB} database rows (32 B), "
                                   f"database sharded x{world}, all-gather + merge", "scaling": "strong"},
            "accepted": int(acc.sum().item()), "gpu_launches_per_step": (m.launch_count() - l0) // steps,
            "compulsory_bytes_per_step": HAM_NDB * 32 + HAM_NQ * 32 * world + HAM_NQ * 13,
            "hbm_gbs": (HAM_NDB * 32 / world) / (ms * 1e-3) / 1e9,
            "note": "bound by integer issue, not by HBM: 4 POPC (16 lanes/clk/SM) + 16 LOP3 + 2.5 VIMNMX + 4 IMAD per pair after carry-save compression of the 8 XOR words"}


def run_gpu(args):
    import torch
    # The C ABI treats a NULL stream as "the handle's own stream"; torch's default stream has handle 0, so all
    # device work of this process runs on one explicit side stream: the kernels, NCCL and the timing events
    # are then ordered on the same stream.
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; liborbx has no CPU fallback")
    torch.cuda.set_device(local)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        return _run_gpu(args, torch)


def _run_gpu(args, torch):
    import torch.distributed as dist
    import orbslam2_b200 as pkg

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; liborbx has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    B = FRAMES_PER_STEP
    ex = pkg.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_width=W, max_height=H, max_batch=B)
    base = base_frames(pkg)
    frames = device_frames(torch, base, B, rank)
    cap = NFEAT + 64
    d_kps = torch.empty((B, cap, 7), dtype=torch.float32, device="cuda")
    d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    assert stream != 0

    # C2: the batch is a sequence of stereo pairs (L R L R ...); Frame::ComputeStereoMatches for every pair runs on the
    # same stream right after the extraction (mb = 0: what the reference's stereo constructor runs with)
    stereo = args.workload == "C2"
    if stereo:
        sm = pkg.ORBmatcher(device=local)
        pairs = B // 2
        d_ur = torch.empty((pairs, cap), dtype=torch.float32, device="cuda")
        d_z = torch.empty((pairs, cap), dtype=torch.float32, device="cuda")
        d_nm = torch.zeros(pairs, dtype=torch.int32, device="cuda")
        mbf = float(np.float32(718.856) * np.float32(0.537))     # KITTI 00-02 fx * baseline

    def step():
        ex.extract_batch_device(frames.data_ptr(), B, W, H, W, W * H, d_kps.data_ptr(), d_desc.data_ptr(), cap,
                                d_n.data_ptr(), stream)
        if stereo:
            sm.stereo_match_batch_device(ex, pairs, 0, 1, 2, d_kps.data_ptr(), d_desc.data_ptr(), d_n.data_ptr(), cap, mbf,
                                         0.0, d_ur.data_ptr(), d_z.data_ptr(), d_nm.data_ptr(), stream)

    # ---- device-resident throughput ("value") ----
    for _ in range(max(args.warmup, 3)):
        step()
    ex.set_profiling(True)
    stage_acc = {}
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = ex.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
        for k, v in ex.last_stage_ms().items():  # CUDA events recorded on the launching stream
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if rank == 0 else None
    launches = ex.launch_count() - launches0
    ex.set_profiling(False)
    value = world * B * args.steps / (ms * 1e-3)
    nkp = int(d_n.sum().item())

    # ---- end to end through the host-buffer C ABI ("e2e") ----
    Be = E2E_FRAMES_PER_STEP
    host = torch.empty((Be, H, W), dtype=torch.uint8).pin_memory()
    host.copy_(frames[:Be].cpu())
    hnp = host.numpy()
    ecap = NFEAT + 64
    h_kps = pkg.pinned_empty((Be, ecap), pkg.KP_DTYPE, device=local)      # results land here without a staging copy
    h_desc = pkg.pinned_empty((Be, ecap, 32), np.uint8, device=local)
    h_n = np.zeros(Be, np.int32)
    e2e_steps = 0 if args.skip_e2e else args.steps
    for _ in range(0 if args.skip_e2e else 3):
        ex.extract_batch_into(hnp, h_kps, h_desc, h_n)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ex.extract_batch_into(hnp, h_kps, h_desc, h_n)   # synchronous: results are in host memory on return
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = world * Be * e2e_steps / e2e_s if e2e_steps else None
    # the same frames as a continuous sequence: orbx_extract_batch_submit / _wait with two steps in flight, each step
    # into its own pinned result arrays; a step's results are waited for before its arrays are handed out again
    stream_fps = None
    if e2e_steps:
        sets = [(h_kps, h_desc, pkg.pinned_empty(Be, np.int32, device=local)),
                (pkg.pinned_empty((Be, ecap), pkg.KP_DTYPE, device=local), pkg.pinned_empty((Be, ecap, 32), np.uint8, device=local),
                 pkg.pinned_empty(Be, np.int32, device=local))]
        def run_stream(k):
            pending = [None, None]
            for i in range(k):
                if pending[i % 2] is not None:
                    ex.extract_batch_wait(pending[i % 2])
                pending[i % 2] = ex.extract_batch_submit(hnp, *sets[i % 2])
            for t in pending:
                if t is not None:
                    ex.extract_batch_wait(t)
        run_stream(3)
        barrier()
        t0 = time.perf_counter()
        run_stream(e2e_steps)
        torch.cuda.synchronize()
        stream_s = max_over_ranks(time.perf_counter() - t0)
        stream_fps = world * Be * e2e_steps / stream_s
        assert np.array_equal(sets[0][2], h_n) and np.array_equal(sets[1][2], h_n), "streaming leg disagrees with the synchronous call"
    rows_per_frame = sum(max(max(f + 3, 4 * 1), 64) + 1 for f in ex.mnFeaturesPerLevel.tolist())
    if e2e_steps:
        assert int(h_n.sum()) > 0 and np.array_equal(h_n, d_n[:Be].cpu().numpy()), "e2e leg disagrees with the device leg"
    e2e_obj = {"value": stream_fps if stream_fps else e2e, "unit": "frames/s", "h2d_bytes_per_step": Be * W * H,
               "d2h_bytes_per_step": Be * rows_per_frame * 60 + Be * 4, "frames_per_step": Be,
               "api": "orbx_extract_batch_submit/_wait (pinned host frames in, pinned host keypoints+descriptors out; the "
                      "steps of a continuous sequence, two in flight, 128-frame chunks rotating over 4 streams; every step's "
                      "H2D and D2H inside the timed region)",
               "synchronous_call": {"value": e2e, "unit": "frames/s",
                                    "api": "orbx_extract_batch, one blocking call per step (64-frame chunks over 4 streams): "
                                           "pays one chunk of pipeline fill and drain per call"}}

    hamming = None if args.skip_hamming else run_hamming(torch, dist, pkg, world, rank, local, args, barrier, max_over_ranks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    stereo_obj = None
    if stereo:
        stereo_obj = {"pairs_per_s": value / 2, "matched_per_pair": float(d_nm.float().mean().item()),
                      "note": "value counts frames; every timed step also runs orbx_stereo_match_batch_device on its "
                              f"{pairs} pairs (3 launches: row table, match + SAD refinement, median filter)"}

    # ---- roofline of the dominant stage, from the event times of the timed region ----
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    if peaks_path.exists():
        peak, peak_src = json.loads(peaks_path.read_text())["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    ab = algorithmic_bytes()
    stages = {}
    for k in ("pyramid", "fast", "octree", "blur", "describe"):
        t = stage_acc[k] / args.steps
        gbs = ab[k] * B / (t * 1e-3) / 1e9 if t > 0 else 0.0
        stages[k] = {"ms_per_step": t, "algorithmic_bytes_per_step": ab[k] * B, "gbs": gbs, "frac_of_hbm": gbs / peak}
    dom = max(("pyramid", "fast", "octree", "blur", "describe"), key=lambda k: stages[k]["ms_per_step"])
    kernel_names = {"pyramid": "resize4_kernel (7 launches)", "fast": "fast_cells_kernel", "octree": "octree_kernel",
                    "blur": "blur_kernel", "describe": "describe_kernel"}
    # DRAM traffic and the limiting pipe of that kernel come from the committed ncu --set full capture (per launch of
    # 512 frames, like `achieved`); null when the digest does not cover this kernel or this launch size
    traffic, limiter = None, None
    digest_path = ROOT / "profiles" / "ncu_digest.json"
    if digest_path.exists():
        dg = json.loads(digest_path.read_text())
        k = dg["kernels"].get(kernel_names[dom])
        if k and dg["frames_per_launch"] == B:
            traffic = k["dram_bytes"]
            limiter = {"pipe": "alu (integer issue)", "pct_of_peak": k.get("alu_pipe_pct"), "issue_pct": k.get("issue_pct"),
                       "source": dg["source"]}
    roof = {"bound": "hbm", "kernel": kernel_names[dom], "achieved": stages[dom]["gbs"], "peak": peak, "unit": "GB/s",
            "frac": stages[dom]["gbs"] / peak, "traffic": traffic, "limiter": limiter, "peak_source": peak_src,
            "share_of_step": stages[dom]["ms_per_step"] / (stage_acc["total"] / args.steps),
            "pipeline_achieved": ab["total"] * value / world / 1e9,
            "pipeline_frac": ab["total"] * value / world / 1e9 / peak,
            "algorithmic_bytes_per_frame": ab["total"]}

    cpu = None
    if world == 1 and not args.skip_cpu:
        cores = os.cpu_count() or 1
        sample_frames = [f for f in frames[:BASE_FRAMES].cpu().numpy()]
        fps, n, med_ms = time_cpu_oracle(sample_frames, 12.0, cores)
        cpu = {"value": fps, "unit": "frames/s", "cores": cores, "kind": cpu_extractor_factory()[1],
               "sample": f"12 s of the same {W}x{H} frames on {cores} threads ({n} frames), one extractor per thread",
               "single_thread_ms_per_frame": med_ms}

    matcher_rows = run_matcher_rows(pkg, local) if (world == 1 and not args.skip_cpu and not args.skip_hamming) else None

    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": B, "input_bytes_per_step_per_gpu": B * W * H,
                       "l2_policy": "inputs larger than L2 (157 MB of distinct frames per step)",
                       "keypoints_per_frame": nkp / B, "parallelism": f"frame sharding x{world}, no collective"},
            "clocks": clocks, "e2e": e2e_obj, "gpu_launches": launches + (3 * args.steps if stereo else 0), "roofline": roof,
            "stages": stages, "stereo": stereo_obj,
            "cpu_baseline": cpu, "hamming": hamming, "matcher_rows": matcher_rows}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
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
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs only: skip the CPU baseline leg")
    ap.add_argument("--skip-hamming", action="store_true", help="profiling runs only: skip the Hamming (config 5) leg")
    ap.add_argument("--workload", default="C1", choices=sorted(WORKLOADS), help="extraction shape (default: the metric's C1)")
    args = ap.parse_args()
    select_workload(args.workload)
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
