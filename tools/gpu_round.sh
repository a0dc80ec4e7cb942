#!/bin/bash
# One gpurun call producing the round's evidence: GPU tests, the full bench line (with the
# CPU baseline), the reference arm, the ncu launch list of the bench command, full ncu
# captures of the sweep kernels (quarter-scale workload: the full-size SW launch overflows
# ncu's counters) and the pipe micro-benchmark.
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
lscpu > gpurun_out/lscpu.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 4 --warmup 3 > gpurun_out/bench_full.log 2>&1
echo "bench exit $?" >> gpurun_out/bench_full.log
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_reference.log 2>&1
echo "reference exit $?" >> gpurun_out/bench_reference.log
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
bash tools/gpu_prof.sh > gpurun_out/prof.log 2>&1
./tools/ubench/pipes > gpurun_out/ubench_pipes.txt 2>&1
python tools/profile_c1.py > gpurun_out/c1.log 2>&1
GRT_METHOD=0 python tools/profile_c1.py > gpurun_out/c1_wavenumber_sweep.log 2>&1
GRT_METHOD=1 python tools/profile_c1.py > gpurun_out/c1_line_sweep.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -c 600 gpurun_out/bench_full.log; tail -c 400 gpurun_out/bench_reference.log; tail -2 gpurun_out/c1.log
