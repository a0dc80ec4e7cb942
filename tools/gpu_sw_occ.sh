#!/bin/bash
# Fused SW solver against resident blocks per SM: throttled with idle shared memory
# (GRT_SW_BLOCKS_PER_SM), then raised with a register cap (-DGRT_SW_MINBLOCKS, rebuild).
mkdir -p gpurun_out
for b in 0 1 2 3 4; do
  if [ $b = 0 ]; then unset GRT_SW_BLOCKS_PER_SM; else export GRT_SW_BLOCKS_PER_SM=$b; fi
  python tools/step_timeline.py 2>&1 | grep "^sw" | grep -o "solver [0-9.]* ms" | sed "s/^/  blocks_per_sm=$b /"
done
unset GRT_SW_BLOCKS_PER_SM
for mb in 6 8 10; do
  make -B -C grtcode_b200/csrc EXTRA="-DGRT_SW_MINBLOCKS=$mb" > gpurun_out/variant_build.log 2>&1 || { echo build failed; continue; }
  grep -A3 "sw_kernelILi64ELb1" grtcode_b200/csrc/solvers.ptxas.log | grep -E "Used|spill" | cut -c1-120
  python tools/step_timeline.py 2>&1 | grep "^sw" | grep -o "solver [0-9.]* ms" | sed "s/^/  minblocks=$mb /"
done
