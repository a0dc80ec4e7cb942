#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_driver.py -m gpu -q 2>&1 | tail -3
echo "=== config-1 column (wing_ns)"
for v in "16 4" "8 4" "16 8"; do
  set -- $v
  GRT_SWEEP_LANES=$1 GRT_SWEEP_PTS=$2 timeout 300 python tools/profile_c1.py 2>&1 | tail -1 | grep -o "wing_ns.: [0-9]*" | sed "s/^/  c1 lanes=$1 pts=$2 /"
done
GRT_SWEEP_LANES=8 GRT_SWEEP_PTS=4 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1
grep '^{' gpurun_out/bench_short.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); r=d.get('rfmip') or {}
    print('  ms/step %.1f %s | rfmip ms/col %s %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step'], r.get('ms_per_column_per_gpu'), r.get('kernel_ms_per_column')))
"
