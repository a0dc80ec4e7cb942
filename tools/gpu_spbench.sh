#!/bin/bash
# The single-precision build's own bench line (beside, never instead of, the f64 headline).
mkdir -p gpurun_out
GRT_PRECISION=single GRT_RFMIP_COLUMNS=16 GRT_RFMIP_CPU_COLUMNS=0 timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sp_bench.log 2>&1
echo "bench exit $?" >> gpurun_out/sp_bench.log
tail -c 2500 gpurun_out/sp_bench.log
