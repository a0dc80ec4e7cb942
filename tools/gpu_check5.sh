#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_driver.py -m gpu -q 2>&1 | tail -3
for m in 0 1; do
  GRT_METHOD=$m timeout 300 python tools/profile_c1.py 2>&1 | tail -1 | cut -c1-320 | sed "s/^/  c1 method=$m /"
done
