#!/bin/bash
# XU-free double->float conversion in the fast class: GRT_F2F_MIX of every four pairs.
mkdir -p gpurun_out
for mix in 0 1 2 4; do
  echo "=== GRT_F2F_MIX=$mix"
  make -B -C grtcode_b200/csrc EXTRA="-DGRT_F2F_MIX=$mix" > gpurun_out/variant_build.log 2>&1 || { echo build failed; tail -5 gpurun_out/variant_build.log; continue; }
  make -C grtcode_b200/host >> gpurun_out/variant_build.log 2>&1
  timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py -m gpu -x -q -k "not full_size" 2>&1 | tail -1
  GRT_BENCH_RFMIP=0 timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('  ms/step %.1f %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step']))
"
  GRT_SWEEP_LANES=16 GRT_SWEEP_PTS=4 timeout 300 python tools/profile_c1.py 2>&1 | tail -1 | grep -o "wing_ns.: [0-9]*" | sed "s/^/  c1 /"
done
