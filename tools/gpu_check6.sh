#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -2
for v in 1 0; do
  GRT_CORE_ORDER=$v timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d.get('rfmip') or {}
        print('  core_order=$v ms/step %.1f %s | rfmip ms/col %s %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step'], r.get('ms_per_column_per_gpu'), r.get('kernel_ms_per_column')))
"
done
