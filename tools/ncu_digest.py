#!/usr/bin/env python3
"""profiles/ncu_digest.json from an `ncu --page raw --csv` export of one pipeline step: per kernel (summed over its
launches) DRAM bytes and pipe utilisation. bench.py copies these into roofline.traffic / roofline.limiter.
Usage: python tools/ncu_digest.py gpurun_out/<tag>/pipeline_full_raw.csv profiles/<name>_pipeline_ncu_raw.csv 512"""
import csv
import hashlib
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def csrc_hash():
    """the same hash bench.py computes over the kernel sources: the digest is only used for the build it describes"""
    h = hashlib.sha256()
    # the sources of the extraction-pipeline kernels (the ones profiles/ncu_digest.json describes)
    for name in ("orb_pattern.inc", "orbx_common.cuh", "orbx_describe.cu", "orbx_fast.cu", "orbx_math.cuh", "orbx_octree.cu",
                 "orbx_pyramid.cu"):
        h.update(name.encode())
        h.update((ROOT / "orb-slam2-_b200" / "csrc" / name).read_bytes())
    return h.hexdigest()[:16]


M = {"dram_rd": "dram__bytes_read.sum", "dram_wr": "dram__bytes_write.sum",
     "alu": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
     "fma": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
     "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active",
     "l1": "l1tex__throughput.avg.pct_of_peak_sustained_active", "ms": "gpu__time_duration.sum"}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
NAMES = {"resize4_kernel": "resize4_kernel (7 launches)"}


def main(path, committed_as, frames):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    acc = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").split("<")[0].strip()
        name = NAMES.get(name, name)
        f = lambda k: float(r[col[M[k]]].replace(",", ""))
        d = acc.setdefault(name, {"dram_bytes": 0.0, "launches": 0, "alu": [], "fma": [], "issue": [], "l1": []})
        d["dram_bytes"] += f("dram_rd") * UNIT[units[col[M["dram_rd"]]]] + f("dram_wr") * UNIT[units[col[M["dram_wr"]]]]
        d["launches"] += 1
        for k in ("alu", "fma", "issue", "l1"):
            d[k].append(f(k))
    out = {"_comment": "per pipeline step (one launch set over frames_per_launch frames), read from an ncu --set full capture; "
                       "bench.py copies these into roofline.traffic / roofline.limiter when csrc_sha16 matches the kernel sources it "
                       "times (run this script with the captured sources checked out). Regenerate with tools/ncu_digest.py.",
           "source": f"profiles/{committed_as} (ncu --set full --clock-control none, B200, {frames} frames per launch)",
           "frames_per_launch": int(frames), "workload": "C1", "csrc_sha16": csrc_hash(), "kernels": {}}
    for name, d in acc.items():
        out["kernels"][name] = {"dram_bytes": round(d["dram_bytes"]), "launches": d["launches"],
                                "alu_pipe_pct": round(max(d["alu"]), 1), "fma_pipe_pct": round(max(d["fma"]), 1),
                                "issue_pct": round(max(d["issue"]), 1), "l1_pct": round(max(d["l1"]), 1)}
    print(json.dumps(out, indent=2))


if __name__ == "__main__":
    main(*sys.argv[1:4])
