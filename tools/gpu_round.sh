#!/bin/bash
# One GPU-box visit: GPU tests, headline bench, the other configurations, the lowered reference models, kernel A/B.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/pytest_gpu.log; tail -n 4 gpurun_out/pytest_gpu.log
python bench.py --steps 30 --warmup 5 2> gpurun_out/bench.err | tee gpurun_out/bench_headline.json | cut -c1-200
val() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['fraction_of_measured_hbm'], d['clocks']['reasons'])"; }
for k in classic ring; do echo "== d3q19 512^3, 26 Globals every iteration, TCLB_KERNEL=$k"; TCLB_KERNEL=$k python bench.py --model d3q19 --shape 512,512,512 --steps 30 --warmup 5 2>> gpurun_out/bench.err | val; done | tee gpurun_out/ab_d3q19.txt
python bench.py --config d3q19 --steps 30 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_d3q19.json | cut -c1-120
python bench.py --config pf_velocity --steps 20 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_pf.json | cut -c1-120
for k in classic ring; do for g in "" "--no-globals"; do echo "== pf_velocity 512^3 TCLB_KERNEL=$k $g"; TCLB_KERNEL=$k python bench.py --config pf_velocity $g --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done; done | tee gpurun_out/ab_pf.txt
for m in ref_d3q27_cumulant ref_d3q27_cumulant_fp ref_d3q19 ref_d3q19_fp d3q19 d3q19_fp d3q27_cumulant d3q27_cumulant_fp; do echo "== $m 512^3 no globals (classic kernel)"; python bench.py --model $m --shape 512,512,512 --no-globals --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done | tee gpurun_out/ref_lowered.txt
