#!/bin/bash
# A/B aid: bench lines (device-resident stages only) of experiment builds made by build.py build_variant.
# Usage (via gpurun): bash tools/variants.sh <tag> name1 name2 ...   ("base" = the in-tree liborbx.so)
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in "$@"; do
  if [ "$v" = base ]; then unset ORBX_LIB; else export ORBX_LIB=$PWD/orb-slam2-_b200/variants/liborbx_$v.so; fi
  timeout 300 python bench.py --steps 10 --warmup 3 --skip-e2e --skip-cpu --skip-hamming --skip-workloads --skip-parity > $OUT/$v.json 2> $OUT/$v.err
  python - "$v" $OUT/$v.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
print(sys.argv[1], "fps %.0f" % d["value"], " ".join("%s %.3f" % (k, v["ms_per_step"]) for k, v in d["stages"].items()))
PY
done
