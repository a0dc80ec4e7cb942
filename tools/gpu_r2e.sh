#!/bin/bash
# Round 2e (two GPUs): multi-device test, binned methods at 0.001 cm-1, N=2 bench with the
# measured-cost rebalancing of the spectral split.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multidevice.py "tests/test_gpu_parity.py::test_binned_methods_at_the_finest_resolution_the_reference_allows" tests/test_gpu_parity.py::test_binned_methods -m gpu -q -s 2>&1 | grep -v "Using " > gpurun_out/r2e_pytest.log
tail -8 gpurun_out/r2e_pytest.log | cut -c1-250
GRT_RFMIP_COLUMNS=8 bash tools/gpu_n2.sh 2 > /dev/null 2>&1
grep '^{' gpurun_out/bench_n2.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['spectral_split']
print('split', s['speedup'], s['efficiency']); [print('  ', p) for p in s['passes']]"
