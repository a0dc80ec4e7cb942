#!/bin/bash
mkdir -p gpurun_out
for mb in 8 10; do
  make -B -C grtcode_b200/csrc EXTRA="-DGRT_SW_MINBLOCKS=$mb" > gpurun_out/variant_build.log 2>&1 || { echo build failed; continue; }
  grep -A3 "sw_kernelILi64ELb1" grtcode_b200/csrc/solvers.ptxas.log | grep -E "Used|spill" | cut -c1-120
  python tools/step_timeline.py 2>&1 | grep "^sw" | grep -o "solver [0-9.]* ms" | sed "s/^/  minblocks=$mb /"
done
