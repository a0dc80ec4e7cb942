#!/bin/bash
# Line-group variants of the wing sweep (GRT_SWEEP_LANES / GRT_SWEEP_PTS, sweep.cuh): the new
# GPU tests with the shipped defaults, parity for every variant, then timings on the config-1
# column, the RFMIP-shaped column and config 2.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for v in "16 4" "8 8" "16 16"; do
  set -- $v
  echo "=== parity LANES=$1 PTS=$2"
  GRT_SWEEP_LANES=$1 GRT_SWEEP_PTS=$2 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
done
echo "=== config-1 column (wing_ns)"
for v in "32 2" "32 4" "16 4" "16 8" "8 8" "8 16"; do
  set -- $v
  GRT_SWEEP_LANES=$1 GRT_SWEEP_PTS=$2 timeout 300 python tools/profile_c1.py 2>&1 | tail -1 | grep -o "wing_ns.: [0-9]*" | sed "s/^/  c1 lanes=$1 pts=$2 /"
done
for m in 0 1; do
  GRT_METHOD=$m timeout 300 python tools/profile_c1.py 2>&1 | tail -1 | cut -c1-300 | sed "s/^/  c1 method=$m /"
done
echo "=== bench (config 2 + RFMIP-shaped columns)"
for v in "32 0" "16 4" "8 8" "16 16" "16 8"; do
  set -- $v
  # PTS=0: the library's own choice per grid
  if [ "$2" = "0" ]; then unset GRT_SWEEP_PTS; else export GRT_SWEEP_PTS=$2; fi
  GRT_SWEEP_LANES=$1 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d.get('rfmip') or {}
        print('  lanes=$1 pts=$2 ms/step %.1f %s | rfmip ms/col %s %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step'], r.get('ms_per_column_per_gpu'), r.get('kernel_ms_per_column')))
"
done
