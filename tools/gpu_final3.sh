#!/bin/bash
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
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; tail -c 300 gpurun_out/bench_full.log; tail -c 200 gpurun_out/bench_reference.log; tail -3 gpurun_out/step_timeline.txt
