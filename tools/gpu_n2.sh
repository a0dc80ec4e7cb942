#!/bin/bash
# Multi-GPU legs under torchrun as the driver launches them (N from $1, default 2).
N=${1:-2}
mkdir -p gpurun_out
GRT_RFMIP_COLUMNS=${GRT_RFMIP_COLUMNS:-225} timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.log 2>&1
echo "bench exit $?" >> gpurun_out/bench_n$N.log
grep -v "Using " gpurun_out/bench_n$N.log | tail -c 3500
