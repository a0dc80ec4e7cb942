#!/bin/bash
# Rebuild the CUDA layer on the GPU box with each flag set of $GRT_FLAGSETS (separated by ';')
# and time config 2 (two steps) and the config-1 column; parity tests for every build.
mkdir -p gpurun_out
IFS=';' read -ra SETS <<< "$GRT_FLAGSETS"
for flags in "${SETS[@]}"; do
  echo "=== $flags"
  make -B -C grtcode_b200/csrc EXTRA="$flags" > gpurun_out/variant_build.log 2>&1 || { echo build failed; continue; }
  make -C grtcode_b200/host >> gpurun_out/variant_build.log 2>&1
  timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
  GRT_BENCH_RFMIP=0 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('  ms/step %.1f %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step']))
"
  for pts in 2 4; do GRT_SWEEP_PTS=$pts timeout 300 python tools/profile_c1.py 2>&1 | tail -1 | grep -o "wing_ns.: [0-9]*" | sed "s/^/  c1 pts=$pts /"; done
done
