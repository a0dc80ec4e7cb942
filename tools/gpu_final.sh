#!/bin/bash
# Final evidence of a round on the shipped build (one gpurun call, ~5 GPU-minutes):
# GPU tests, smoke, bench line (with CPU baseline), reference arm, step timeline, ncu launch
# list of the bench command, config-1 column for the three methods, and ncu full captures of
# the coarse-grid wing kernel (config-1 column) and of the fused SW solver.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py --steps 4 --warmup 3 > gpurun_out/bench_full.log 2>&1
echo "bench exit $?" >> gpurun_out/bench_full.log
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_reference.log 2>&1
echo "reference exit $?" >> gpurun_out/bench_reference.log
python tools/step_timeline.py > gpurun_out/step_timeline.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python tools/profile_c1.py > gpurun_out/c1.log 2>&1
GRT_METHOD=0 python tools/profile_c1.py > gpurun_out/c1_wavenumber_sweep.log 2>&1
GRT_METHOD=1 python tools/profile_c1.py > gpurun_out/c1_line_sweep.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:wing_sweep -s 2 -c 1 \
    -f -o gpurun_out/prof_wing_c1 python tools/profile_c1.py > gpurun_out/ncu_wing_c1.log 2>&1
GRT_BENCH_RFMIP=0 ncu --set full --clock-control none --import-source on -k regex:sw_kernel -s 1 -c 1 \
    -f -o gpurun_out/prof_sw_solver python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_sw_solver.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; tail -c 400 gpurun_out/bench_full.log; tail -c 300 gpurun_out/bench_reference.log; tail -3 gpurun_out/step_timeline.txt
