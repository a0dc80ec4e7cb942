#!/bin/bash
# The single-precision build (GRT_PRECISION=single): parity against the reference's own
# -DSINGLE_PRECISION build, then its bench line (beside, never instead of, the f64 headline).
mkdir -p gpurun_out
GRT_PRECISION=single timeout 600 python -m tests.sp_check 2>&1 | grep SPCHECK > gpurun_out/sp_check.txt
timeout 900 python -m pytest tests/test_gpu_single.py -m gpu -x -q > gpurun_out/sp_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/sp_pytest.log; tail -3 gpurun_out/sp_pytest.log
GRT_PRECISION=single GRT_RFMIP_COLUMNS=16 GRT_RFMIP_CPU_COLUMNS=0 timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sp_bench.log 2>&1
echo "bench exit $?"; grep '^{' gpurun_out/sp_bench.log | tail -1 > gpurun_out/sp_bench.json; tail -c 600 gpurun_out/sp_bench.json
