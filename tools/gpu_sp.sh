#!/bin/bash
mkdir -p gpurun_out
GRT_PRECISION=single timeout 600 python -m tests.sp_check > gpurun_out/sp_check.log 2>&1
echo "exit $?" >> gpurun_out/sp_check.log
grep "SPCHECK\|exit\|Error\|error" gpurun_out/sp_check.log | cut -c1-1800 | tail -20
timeout 900 python -m pytest tests/test_gpu_solvers.py tests/test_gpu_parity.py tests/test_gpu_single.py -m gpu -x -q > gpurun_out/sp_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/sp_pytest.log
tail -5 gpurun_out/sp_pytest.log
