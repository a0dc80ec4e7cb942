#!/bin/bash
mkdir -p gpurun_out
GRT_METHOD=0 GRT_REPS=3 python tools/profile_c1.py > gpurun_out/c1_m0.log 2>&1 &&
GRT_METHOD=0 GRT_REPS=3 ncu --set full --clock-control none --import-source on -k regex:bin_gather -s 2 -c 1 \
    -f -o gpurun_out/prof_bin_gather python tools/profile_c1.py > gpurun_out/ncu_bin.log 2>&1
tail -2 gpurun_out/ncu_bin.log
