#!/bin/bash
mkdir -p gpurun_out
python tools/ab_run.py $GRT_AB_VARIANTS 2>&1 | tee gpurun_out/ab_${GRT_AB_TAG:-run}.txt
