#!/bin/bash
# Final evidence of the round on the shipped build: GPU tests, bench line (with CPU baseline),
# reference arm, ncu launch list of the bench command, config-1 column for the three methods.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py --steps 4 --warmup 3 > gpurun_out/bench_full.log 2>&1
echo "bench exit $?" >> gpurun_out/bench_full.log
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_reference.log 2>&1
echo "reference exit $?" >> gpurun_out/bench_reference.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python tools/profile_c1.py > gpurun_out/c1.log 2>&1
GRT_METHOD=0 python tools/profile_c1.py > gpurun_out/c1_wavenumber_sweep.log 2>&1
GRT_METHOD=1 python tools/profile_c1.py > gpurun_out/c1_line_sweep.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; tail -c 400 gpurun_out/bench_full.log; tail -c 300 gpurun_out/bench_reference.log; tail -1 gpurun_out/c1.log
