#!/bin/bash
# A/B aid for the brute-force matcher: config 5 through bench.py for experiment builds (build.py build_variant).
# Usage (via gpurun): bash tools/ham_variants.sh <tag> name1 name2 ...   ("base" = the in-tree liborbx.so)
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in "$@"; do
  if [ "$v" = base ]; then unset ORBX_LIB; else export ORBX_LIB=$PWD/orb-slam2-_b200/variants/liborbx_$v.so; fi
  timeout 300 python bench.py --steps 5 --warmup 3 --skip-e2e --skip-cpu --skip-workloads > $OUT/$v.json 2> $OUT/$v.err
  python - "$v" $OUT/$v.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
h = d["hamming"]
print(sys.argv[1], "Gpairs/s %.1f" % h["value"], "ms %.2f" % h["ms_per_step"], h["parity"]["result"], "alu frac %.3f" % h["roofline"]["frac"])
PY
done
