#!/bin/bash
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["dtype"], d["config"]["lattice"], d["value"], d["ms_per_step"], d["roofline"]["bytes_per_update"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"])'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -4 gpurun_out/pytest_gpu.log
echo "== headline"; timeout 900 python bench.py --steps 30 --warmup 5 2>gpurun_out/h.err | tee gpurun_out/bench_n1.json | python -c "$P"; tail -2 gpurun_out/h.err
echo "== cumulant half-shift storage"; timeout 900 python bench.py --steps 20 --warmup 5 --precision float_halfshift --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
echo "== d3q19 half-shift storage"; timeout 900 python bench.py --steps 20 --warmup 5 --model d3q19 --shape 512,512,512 --precision float_halfshift --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
echo "== d2q9 half-shift storage"; timeout 900 python bench.py --steps 20 --warmup 5 --model d2q9 --shape 8192,8192,1 --precision float_halfshift --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
echo "== d2q9 float"; timeout 900 python bench.py --steps 20 --warmup 5 --model d2q9 --shape 8192,8192,1 --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
echo "== pf 512^3"; timeout 900 python bench.py --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 512,512,512 --no-globals 2>gpurun_out/pf.err | python -c "$P"; tail -5 gpurun_out/pf.err
