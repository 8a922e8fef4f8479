#!/bin/bash
# One GPU-box visit: GPU tests, headline bench, the other configurations, kernel A/B, one ncu capture (a window of kernels).
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log | tail -8
python bench.py --steps 30 --warmup 5 2> gpurun_out/bench.err | tee gpurun_out/bench_headline.json | cut -c1-400
LIBS="${LIBS:-d3q27_cumulant_pb4 d3q27_cumulant_pb5}" bash tools/ab_persist.sh > /dev/null 2>&1; cat gpurun_out/ab_kernel2.txt
for k in classic ring; do echo "== d3q19 globals every iteration, TCLB_KERNEL=$k"; TCLB_KERNEL=$k python bench.py --model d3q19 --shape 512,512,512 --steps 30 --warmup 5 2>> gpurun_out/bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['fraction_of_measured_hbm'])"; done | tee gpurun_out/ab_d3q19.txt
python bench.py --config d3q19 --steps 30 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_d3q19.json | cut -c1-300
python bench.py --config pf_velocity --steps 20 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_pf.json | cut -c1-300
for m in ref_d3q27_cumulant ref_d3q19; do echo "== $m no globals"; python bench.py --model $m --shape 512,512,512 --no-globals --steps 20 --warmup 5 2>> gpurun_out/bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['fraction_of_measured_hbm'])"; done | tee gpurun_out/ref_lowered.txt
python tools/ncu_window.py > gpurun_out/window_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/prof_window python tools/ncu_window.py > gpurun_out/ncu_window.log 2>&1
tail -n 3 gpurun_out/window_plain.log; tail -n 3 gpurun_out/ncu_window.log
ls -la gpurun_out/*.ncu-rep
