#!/bin/bash
# GPU tests with the shipped build, then the per-class cycle timers (debug rebuild on the box).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|^FAILED|^ERROR" gpurun_out/pytest_gpu.log | head -40
make -B -C grtcode_b200/csrc EXTRA=-DGRT_DBG_TIMERS=1 > gpurun_out/timers_build.log 2>&1 && make -C grtcode_b200/host >> gpurun_out/timers_build.log 2>&1
for dw in 0.1 0.01; do
  timeout 600 python tools/class_timers.py --dw $dw > gpurun_out/class_timers_$dw.txt 2>&1
  cat gpurun_out/class_timers_$dw.txt
done
