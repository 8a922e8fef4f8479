#!/bin/bash
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["config"]["lattice"], d["value"], d["ms_per_step"], "graph_iters", d.get("graph_iterations"), "launches", d["gpu_launches"])'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -6 gpurun_out/pytest_gpu.log
echo "== headline"; timeout 900 python bench.py --steps 30 --warmup 5 2>gpurun_out/h.err | tee gpurun_out/bench_n1.json | python -c "$P"; tail -2 gpurun_out/h.err
for shape in 1024,128,1 256,256,1 4096,1024,1; do
echo "== d2q9 $shape graph"; timeout 300 python bench.py --steps 2000 --warmup 100 --model d2q9 --shape $shape --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
echo "== d2q9 $shape plain"; timeout 300 python bench.py --steps 2000 --warmup 100 --model d2q9 --shape $shape --no-globals --no-graph 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
done
for shape in 64,64,64 128,128,128; do
echo "== cumulant $shape graph"; timeout 300 python bench.py --steps 1000 --warmup 50 --shape $shape --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
echo "== cumulant $shape plain"; timeout 300 python bench.py --steps 1000 --warmup 50 --shape $shape --no-globals --no-graph 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
done
echo "== cumulant headline size noglob (graph path)"; timeout 900 python bench.py --steps 30 --warmup 10 --no-globals 2>gpurun_out/h.err | python -c "$P"; tail -2 gpurun_out/h.err
