#!/bin/bash
# Staged shortwave solver: parity, A/B against the thread-per-wavenumber kernel, DRAM bytes.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_solvers.py tests/test_gpu_parity.py tests/test_c_dropin.py tests/test_driver.py tests/test_gpu_headline.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2j_pytest.log; tail -4 gpurun_out/r2j_pytest.log
for k in staged thread; do
  GRT_SW_KERNEL=$k GRT_RFMIP_COLUMNS=8 GRT_RFMIP_CPU_COLUMNS=0 timeout 900 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2j_bench_$k.log 2>&1
  echo "bench $k exit $?"; tail -1 gpurun_out/r2j_bench_$k.log > gpurun_out/r2j_bench_$k.json
done
python - <<'PY'
import json
for n in ("staged","thread"):
    try:
        d=json.loads(open(f"gpurun_out/r2j_bench_{n}.json").read())
        print(n, "ms/step", d["ms_per_step"], "rfmip col/s", d["rfmip"]["columns_per_s"])
    except Exception as e: print(n, "ERR", e)
PY
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct -k regex:"sw_staged_kernel|sw_kernel" --clock-control none --csv --log-file gpurun_out/r2j_sw_dram.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2j_ncu.log 2>&1
echo "ncu exit $?"; grep -c sw_staged gpurun_out/r2j_sw_dram.csv; tail -8 gpurun_out/r2j_sw_dram.csv | cut -c1-300
