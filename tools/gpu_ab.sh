#!/bin/bash
# A/B of a compile-time sweep option on the GPU box: bench with the shipped build, rebuild
# the CUDA layer with $GRT_AB_FLAGS, run the arithmetic selftest + parity + bench again.
mkdir -p gpurun_out
run() {
  timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_$1.log 2>&1
  echo "$1 exit $?"
  grep '^{' gpurun_out/bench_$1.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); r=d['roofline']
    print('  ms/step %.1f samples/s %.3e %s' % (d['ms_per_step'], r['samples_per_s'], r.get('kernel_ms_per_step')))
"
}
run base
make -B -C grtcode_b200/csrc EXTRA="$GRT_AB_FLAGS" > gpurun_out/ab_build.log 2>&1 && make -C grtcode_b200/host >> gpurun_out/ab_build.log 2>&1
echo "rebuild exit $?"
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "
from grtcode_b200 import api
api.load()
print('selftest mismatches over 2^32 operands:', api.selftest_arithmetic(1<<32, seed=777))
"
run variant
