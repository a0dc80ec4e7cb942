#!/bin/bash
# Round-2 third GPU call: the allocation-free entry point's test alone in a fresh process
# (double, then the single-precision build), then the whole GPU suite with it last.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s > gpurun_out/r2c_noalloc_alone.log 2>&1
echo "alone exit $?" >> gpurun_out/r2c_noalloc_alone.log
GRT_PRECISION=single timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s > gpurun_out/r2c_noalloc_alone_sp.log 2>&1
echo "alone single exit $?" >> gpurun_out/r2c_noalloc_alone_sp.log
grep -v "Using \|^$" gpurun_out/r2c_noalloc_alone.log | tail -8
grep -v "Using \|^$" gpurun_out/r2c_noalloc_alone_sp.log | tail -8
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2c_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2c_pytest_gpu.log
grep -v "Using \|^$\|checked build defined" gpurun_out/r2c_pytest_gpu.log | tail -12
