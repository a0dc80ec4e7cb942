#!/bin/bash
# Single build with float partial sums in every line class: parity, then both bench lines.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py tests/test_gpu_parity.py tests/test_gpu_headline.py tests/test_gpu_kernels.py -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2i_pytest.log; tail -4 gpurun_out/r2i_pytest.log
GRT_PRECISION=single GRT_RFMIP_COLUMNS=16 GRT_RFMIP_CPU_COLUMNS=0 timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2i_sp_bench.log 2>&1
echo "sp bench exit $?"; tail -1 gpurun_out/r2i_sp_bench.log > gpurun_out/r2i_sp_bench.json
GRT_RFMIP_COLUMNS=16 GRT_RFMIP_CPU_COLUMNS=0 timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2i_dp_bench.log 2>&1
echo "dp bench exit $?"; tail -1 gpurun_out/r2i_dp_bench.log > gpurun_out/r2i_dp_bench.json
python - <<'PY'
import json
for n in ("sp","dp"):
    try:
        d=json.loads(open(f"gpurun_out/r2i_{n}_bench.json").read())
        print(n, "ms/step", d["ms_per_step"], "rfmip", d.get("rfmip",{}).get("ms_per_column"), d.get("kernel_ms_per_step"))
    except Exception as e: print(n, "ERR", e)
PY
