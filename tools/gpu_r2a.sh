#!/bin/bash
# Round-2 first GPU call: the whole GPU suite (new headline / K1 tests included) and a short bench.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2a_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2a_pytest_gpu.log
grep -v "Using \|^$" gpurun_out/r2a_pytest_gpu.log | tail -80
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r2a_bench.log 2>&1
echo "bench exit $?" >> gpurun_out/r2a_bench.log
tail -c 1500 gpurun_out/r2a_bench.log
