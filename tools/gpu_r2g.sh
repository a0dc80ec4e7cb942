#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py tests/test_gpu_headline.py tests/test_driver.py tests/test_c_dropin.py -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2g_pytest.log; tail -4 gpurun_out/r2g_pytest.log
python tools/setup_profile.py 2>&1 | grep -v "Using " | tee gpurun_out/r2g_setup.txt
