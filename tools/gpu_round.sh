#!/bin/bash
# One gpurun call: GPU tests, the config-1 timing, the bench line, then the ncu
# launch list and one full capture of the dominant kernel (same bench command).
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
lscpu > gpurun_out/lscpu.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python tools/profile_c1.py > gpurun_out/c1.log 2>&1
python bench.py --steps 2 --warmup 3 > gpurun_out/bench_full.log 2>&1
echo "bench exit $?" >> gpurun_out/bench_full.log
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$BENCH > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
$BENCH > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:wing_sweep -s 6 -c 2 \
    -f -o gpurun_out/prof_wing $BENCH > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -5 gpurun_out/c1.log; tail -2 gpurun_out/bench_full.log
