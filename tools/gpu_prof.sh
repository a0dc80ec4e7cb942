#!/bin/bash
# ncu full capture of the sweep kernels on a quarter-scale config 2 (shorter launches:
# the 470 ms SW launch of the full workload overflows ncu's counters).
mkdir -p gpurun_out
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --scale 0.25"
$BENCH > gpurun_out/plain_q.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:wing_sweep -s 2 -c 2 \
    -f -o gpurun_out/prof_sweep_q $BENCH > gpurun_out/ncu_q.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:core_sweep -s 3 -c 1 \
    -f -o gpurun_out/prof_core_q $BENCH > gpurun_out/ncu_core_q.log 2>&1
# full-size longwave launch (the first wing launch of the run): DRAM traffic of a real launch
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
GRT_BENCH_RFMIP=0 ncu --set full --clock-control none --import-source on -k regex:wing_sweep -s 0 -c 1 \
    -f -o gpurun_out/prof_wing_lw_full $FULL > gpurun_out/ncu_lw_full.log 2>&1
# DRAM bytes of both full-size launches (two metrics only: one replay pass, no counter overflow)
GRT_BENCH_RFMIP=0 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
    --clock-control none -k regex:wing_sweep -s 2 -c 2 $FULL > gpurun_out/ncu_wing_dram_full.log 2>&1
grep -E "wing_sweep|dram__bytes|gpu__time" gpurun_out/ncu_wing_dram_full.log | tail -8
tail -2 gpurun_out/ncu_q.log; tail -2 gpurun_out/ncu_core_q.log; tail -2 gpurun_out/ncu_lw_full.log
grep '^{' gpurun_out/plain_q.log | cut -c1-400
