#!/bin/bash
# Quick GPU check of the shipped build: every GPU test, the config-1 column, one short bench.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|^FAILED|^ERROR" gpurun_out/pytest_gpu.log | head -40
for m in 2 0 1; do
  GRT_METHOD=$m timeout 300 python tools/profile_c1.py 2>&1 | tail -1 | cut -c1-320 | sed "s/^/  c1 method=$m /"
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1
grep '^{' gpurun_out/bench_short.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); r=d.get('rfmip') or {}
    print('  ms/step %.1f %s | rfmip ms/col %s %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step'], r.get('ms_per_column_per_gpu'), r.get('kernel_ms_per_column')))
"
