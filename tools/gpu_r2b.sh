#!/bin/bash
# Round-2 second GPU call: the reworked bench (all legs, short), the reference arm, and the
# un-windowed reference run of the whole LW band.
mkdir -p gpurun_out
nproc > gpurun_out/r2b_nproc.txt; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/r2b_nproc.txt
GRT_RFMIP_COLUMNS=${GRT_RFMIP_COLUMNS:-16} timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r2b_bench.log 2>&1
echo "bench exit $?" >> gpurun_out/r2b_bench.log
tail -c 3000 gpurun_out/r2b_bench.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2b_reference.log 2>&1
echo "reference exit $?" >> gpurun_out/r2b_reference.log
tail -c 1200 gpurun_out/r2b_reference.log
timeout 900 python tools/reference_full_band.py lw > gpurun_out/r2b_full_band.log 2>&1
echo "full band exit $?" >> gpurun_out/r2b_full_band.log
tail -c 1500 gpurun_out/r2b_full_band.log
