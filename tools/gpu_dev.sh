#!/bin/bash
# Development round on the GPU box: parity tests, then the bench under sweep variants.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for v in ${GRT_VARIANTS:-8:1 8:2 4:1}; do
  set -- ${v/:/ }
  GRT_SWEEP_PTS=$1 GRT_SWEEP_NL=$2 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_pts$1_nl$2.log 2>&1
  echo "PTS=$1 NL=$2 exit $?"
  grep '^{' gpurun_out/bench_pts$1_nl$2.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); r=d['roofline']
    print('  ms/step %.1f value %.3e samples/s %.3e shares %s' % (d['ms_per_step'], d['value'], r['samples_per_s'], r.get('kernel_ms_per_step', r['kernel_share_of_step'])))
"
done
for v in ${GRT_C1_VARIANTS:-4:1 2:1}; do
  set -- ${v/:/ }
  GRT_SWEEP_PTS=$1 GRT_SWEEP_NL=$2 timeout 300 python tools/profile_c1.py 2>&1 | tail -1
done
