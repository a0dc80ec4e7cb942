#!/bin/bash
mkdir -p gpurun_out
GRT_PRECISION=single timeout 600 python tools/sp_debug.py 2>&1 | grep -v "Using " > gpurun_out/sp_debug.log
tail -70 gpurun_out/sp_debug.log | cut -c1-260
