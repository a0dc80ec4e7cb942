#!/bin/bash
# Last evidence of the round: ncu full captures of the coarse-grid wing kernel (config-1
# column) and of the fused SW solver, then the bench line, the reference arm and the launch list.
mkdir -p gpurun_out
python tools/profile_c1.py > gpurun_out/c1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:wing_sweep -s 2 -c 1 \
    -f -o gpurun_out/prof_wing_c1 python tools/profile_c1.py > gpurun_out/ncu_wing_c1.log 2>&1
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
GRT_BENCH_RFMIP=0 ncu --set full --clock-control none --import-source on -k regex:sw_kernel -s 1 -c 1 \
    -f -o gpurun_out/prof_sw_solver $BENCH > gpurun_out/ncu_sw_solver.log 2>&1
python bench.py --steps 4 --warmup 3 > gpurun_out/bench_full.log 2>&1
echo "bench exit $?" >> gpurun_out/bench_full.log
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_reference.log 2>&1
echo "reference exit $?" >> gpurun_out/bench_reference.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_wing_c1.log; tail -2 gpurun_out/ncu_sw_solver.log; tail -c 300 gpurun_out/bench_full.log; tail -c 200 gpurun_out/bench_reference.log
