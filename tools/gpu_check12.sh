#!/bin/bash
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["config"]["lattice"], d["value"], d["ms_per_step"], "graph_iters", d.get("graph_iterations"))'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -4 gpurun_out/pytest_gpu.log
for rep in 1 2; do
echo "== cumulant full noglob graph"; timeout 900 python bench.py --steps 30 --warmup 10 --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
echo "== cumulant full noglob plain"; timeout 900 python bench.py --steps 30 --warmup 10 --no-globals --no-graph 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
done
echo "== d3q19 512 graph"; timeout 900 python bench.py --steps 30 --warmup 10 --model d3q19 --shape 512,512,512 --no-globals 2>gpurun_out/h.err | python -c "$P"
echo "== d3q19 512 plain"; timeout 900 python bench.py --steps 30 --warmup 10 --model d3q19 --shape 512,512,512 --no-globals --no-graph 2>gpurun_out/h.err | python -c "$P"
