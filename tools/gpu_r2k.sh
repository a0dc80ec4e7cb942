#!/bin/bash
# Core gather (no atomics) against the scatter kernel: parity, then both bench lines.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py tests/test_gpu_headline.py tests/test_gpu_k1.py -m gpu -x -q -s > gpurun_out/r2k_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2k_pytest.log; grep "gather vs scatter\|passed\|failed\|Error\|exit" gpurun_out/r2k_pytest.log | tail -12
for k in gather scatter; do
  GRT_CORE=$k GRT_RFMIP_COLUMNS=8 GRT_RFMIP_CPU_COLUMNS=0 timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_bench_$k.log 2>&1
  echo "bench $k exit $?"; tail -1 gpurun_out/r2k_bench_$k.log > gpurun_out/r2k_bench_$k.json
done
python - <<'PY'
import json
for n in ("gather","scatter"):
    try:
        d=json.loads(open(f"gpurun_out/r2k_bench_{n}.json").read())
        print(n, "ms/step", round(d["ms_per_step"],2), d["roofline"]["kernel_ms_per_step"], "rfmip", d["rfmip"]["kernel_ms_per_column"])
    except Exception as e: print(n, "ERR", e)
PY
