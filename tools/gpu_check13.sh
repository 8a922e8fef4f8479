#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --shape 512,256,256 --no-globals --no-graph"
$CMD > gpurun_out/plain_c.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 3 -c 1 -o gpurun_out/prof_cumulant10_noglob $CMD > gpurun_out/ncu_full2.log 2>&1
tail -2 gpurun_out/ncu_full2.log
