#!/bin/bash
# ncu --set full capture of ONE launch of one kernel of the pipeline (after warm-up), with source counters.
# Usage: bash tools/ncu_one.sh <tag> <kernel-regex> [launch-skip]
TAG=$1; KRE=$2; SKIP=${3:-3}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python bench.py --steps 1 --warmup 3 --skip-e2e --skip-cpu --skip-hamming --skip-workloads --skip-parity"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$KRE" --launch-skip $SKIP --launch-count 1 \
  -o $OUT/one -f $CMD > $OUT/ncu_one.log 2>&1
echo "ncu exit $?"
ncu -i $OUT/one.ncu-rep --page raw --csv > $OUT/one_raw.csv 2>/dev/null
ncu -i $OUT/one.ncu-rep --page details > $OUT/one_details.txt 2>/dev/null
ncu -i $OUT/one.ncu-rep --page source --csv > $OUT/one_source.csv 2>/dev/null
rm -f $OUT/one.ncu-rep   # gpurun only brings back 64 MiB
python tools/ncu_summary.py $OUT/one_raw.csv
python tools/sass_regions.py $OUT/one_source.csv 1.0 > $OUT/one_sass_regions.txt 2>&1
gzip -f $OUT/one_source.csv
