#!/bin/bash
# Timing experiments: rebuild the CUDA layer with each flag set in $GRT_EXP (';'-separated)
# and run the bench (no parity: some flags deliberately break the arithmetic).
mkdir -p gpurun_out
IFS=';' read -ra SETS <<< "$GRT_EXP"
for i in "${!SETS[@]}"; do
  make -B -C grtcode_b200/csrc EXTRA="${SETS[$i]}" > gpurun_out/exp_build_$i.log 2>&1 && make -C grtcode_b200/host >> gpurun_out/exp_build_$i.log 2>&1 || { echo "build $i failed"; tail -5 gpurun_out/exp_build_$i.log; continue; }
  timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline ${GRT_EXP_ARGS} > gpurun_out/bench_exp_$i.log 2>&1
  echo "[${SETS[$i]}] exit $?"
  grep '^{' gpurun_out/bench_exp_$i.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); r=d['roofline']
    print('  ms/step %.1f samples/s %.3e %s' % (d['ms_per_step'], r['samples_per_s'], r.get('kernel_ms_per_step')))
"
done
