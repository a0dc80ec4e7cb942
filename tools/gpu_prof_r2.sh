#!/bin/bash
# Round-2 profiling evidence (one GPU): ncu launch list of the bench command, DRAM bytes of
# the wing launches at full size, --set full captures of the coarse-grid wing kernel (config-1
# column), the fused SW solver and K1; each only after the plain command exited 0.
mkdir -p gpurun_out
export GRT_RFMIP_COLUMNS=4 GRT_RFMIP_CPU_COLUMNS=0
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/p2_bench_plain.log 2>&1 || { echo "plain bench failed"; tail -5 gpurun_out/p2_bench_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/p2_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/p2_ncu_launches.log 2>&1
echo "launch list exit $?"
GRT_BENCH_RFMIP=0 GRT_BENCH_REFAPI=0 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:wing_sweep -s 2 -c 2 --csv --log-file gpurun_out/p2_wing_dram.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/p2_ncu_dram.log 2>&1
echo "wing dram exit $?"
python tools/profile_c1.py > gpurun_out/p2_c1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:wing_sweep -s 2 -c 1 \
    -f -o gpurun_out/p2_wing_c1 python tools/profile_c1.py > gpurun_out/p2_ncu_wing_c1.log 2>&1
echo "wing c1 full exit $?"
GRT_BENCH_RFMIP=0 GRT_BENCH_REFAPI=0 ncu --set full --clock-control none --import-source on -k regex:sw_kernel -s 1 -c 1 \
    -f -o gpurun_out/p2_sw_solver python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/p2_ncu_sw.log 2>&1
echo "sw solver full exit $?"
GRT_BENCH_RFMIP=0 GRT_BENCH_REFAPI=0 ncu --set full --clock-control none --import-source on -k regex:line_prep -s 3 -c 1 \
    -f -o gpurun_out/p2_line_prep python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/p2_ncu_k1.log 2>&1
echo "K1 full exit $?"
ls -la gpurun_out/*.ncu-rep
