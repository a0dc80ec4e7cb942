#!/bin/bash
# Round 2d: solver tests after the beam-sharing change, A/B bench, config-5 slab run.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_solvers.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2d_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2d_pytest.log; tail -3 gpurun_out/r2d_pytest.log
python tools/ab_run.py base 2>&1 | tee gpurun_out/r2d_ab.txt
GRT_C5_COLUMNS=${GRT_C5_COLUMNS:-512} timeout 900 python tools/config5_run.py 2> gpurun_out/config5.err | grep '^{' > gpurun_out/config5.json
cat gpurun_out/config5.json | cut -c1-1500
