#!/bin/bash
mkdir -p gpurun_out
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --scale 0.25"
GRT_BENCH_RFMIP=0 $BENCH > gpurun_out/plain_q.log 2>&1 &&
GRT_BENCH_RFMIP=0 ncu --set full --clock-control none --import-source on -k regex:core_sweep -s 3 -c 1 \
    -f -o gpurun_out/prof_core_q $BENCH > gpurun_out/ncu_core_q.log 2>&1
tail -2 gpurun_out/ncu_core_q.log
