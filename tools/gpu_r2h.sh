#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_setup.py tests/test_gpu_parity.py tests/test_c_dropin.py tests/test_driver.py -m gpu -x -q -s > gpurun_out/r2h_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2h_pytest.log; grep -v "^Using\|^$" gpurun_out/r2h_pytest.log | tail -25
python tools/setup_profile.py 2>&1 | grep -v "Using " | tee gpurun_out/r2h_setup.txt
