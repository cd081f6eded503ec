#!/bin/bash
# One GPU-box cycle: parity tests, bench line, ncu launch list, ncu --set full of the pipeline kernels.
# Usage (from the repo root, via gpurun): bash tools/gpu_cycle.sh <tag> [quick]
TAG=${1:-r1}
MODE=${2:-full}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1
echo "pytest exit $?" >> $OUT/pytest.log
tail -15 $OUT/pytest.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err
echo "bench exit $?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
[ "$MODE" = quick ] && exit 0
CMD="python bench.py --steps 1 --warmup 3 --skip-e2e --skip-cpu --skip-hamming --skip-workloads --skip-parity"
KRE='fast_cells|describe_kernel|blur_kernel|resize4|resize_kernel|octree_kernel'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$KRE" --launch-skip 33 -c 11 --csv \
  --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:"$KRE" --launch-skip 33 --launch-count 11 \
  -o $OUT/pipeline_full -f $CMD > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i $OUT/pipeline_full.ncu-rep --page raw --csv > $OUT/pipeline_full_raw.csv 2>/dev/null
rm -f $OUT/pipeline_full.ncu-rep   # gpurun only brings back 64 MiB
# per-line instruction / stall counters (source page) of the two heaviest kernels, one launch each
bash tools/ncu_one.sh $TAG/fast fast_cells 3 > $OUT/ncu_one_fast.log 2>&1
bash tools/ncu_one.sh $TAG/describe describe_kernel 3 > $OUT/ncu_one_describe.log 2>&1
# matcher / stereo kernels: one small scripted pass (tools/match_probe.py), launch list + full capture
MCMD="python tools/match_probe.py"
timeout 300 $MCMD > $OUT/match_probe.log 2>&1
echo "match probe exit $?"; tail -5 $OUT/match_probe.log
timeout 600 ncu --set full --clock-control none --import-source on \
  -k regex:'bf2_kernel|window_kernel|grid_build|stereo_match|stereo_median|merge_kernel|block_distance|distinctive_kernel|undistort_keys|voc_descend' --launch-count 40 \
  -o $OUT/match_full -f $MCMD > $OUT/ncu_match.log 2>&1
echo "ncu match exit $?"
ncu -i $OUT/match_full.ncu-rep --page raw --csv > $OUT/match_full_raw.csv 2>/dev/null
rm -f $OUT/match_full.ncu-rep
# the key-frame / frame-epilogue / vocabulary kernels come last in the probe: a capture of their own
timeout 600 ncu --set full --clock-control none --import-source on \
  -k regex:'block_distance|distinctive_kernel|undistort_keys|voc_descend|stereo_' --launch-count 14 \
  -o $OUT/rows_full -f $MCMD > $OUT/ncu_rows.log 2>&1
echo "ncu rows exit $?"
ncu -i $OUT/rows_full.ncu-rep --page raw --csv > $OUT/rows_full_raw.csv 2>/dev/null
rm -f $OUT/rows_full.ncu-rep
ls -la $OUT
