#!/bin/bash
# ncu captures of the hot kernels of the models lowered from the reference's files (after the C() block rewriting)
mkdir -p gpurun_out
C1="python bench.py --steps 3 --warmup 3 --model ref_d3q19 --shape 512,512,512 --no-globals --no-graph"
$C1 > gpurun_out/ref19_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 3 -c 1 -o gpurun_out/prof_ref_d3q19 $C1 > gpurun_out/ncu_ref19.log 2>&1; tail -1 gpurun_out/ncu_ref19.log
C2="python bench.py --steps 2 --warmup 3 --model ref_d3q27_pf_velocity --shape 256,256,256 --no-globals --no-graph"
$C2 > gpurun_out/pf_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 12 -c 4 -o gpurun_out/prof_ref_pf $C2 > gpurun_out/ncu_pf.log 2>&1; tail -1 gpurun_out/ncu_pf.log
