#!/bin/bash
# One GPU-box visit after a rebuild: the whole GPU suite (-s for the timings some tests print), headline, the other configurations.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s 2>&1 | grep -v "arena\]" | tail -30 > gpurun_out/pytest_gpu.log; tail -n 12 gpurun_out/pytest_gpu.log
python bench.py --steps 30 --warmup 5 2> gpurun_out/bench.err | tee gpurun_out/bench_headline.json | cut -c1-160
python bench.py --config d3q19 --steps 30 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_d3q19.json | cut -c1-120
python bench.py --config pf_velocity --steps 20 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_pf.json | cut -c1-120
