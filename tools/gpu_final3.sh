#!/bin/bash
# round-end check after the adjoint work: GPU tests, smoke, default bench, adjoint throughput
mkdir -p gpurun_out
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== headline"; timeout 900 python bench.py > gpurun_out/bench_default.json 2>gpurun_out/h.err; echo "rc=$?"; cat gpurun_out/bench_default.json | cut -c1-600
echo "== reference arm"; python bench.py --impl reference | cut -c1-200
echo "== adjoint throughput"
timeout 600 python tools/adjoint_bench.py d2q9 float 2048 2048 1 20 2>&1 | tail -1
timeout 600 python tools/adjoint_bench.py d2q9 double 2048 2048 1 20 2>&1 | tail -1
timeout 600 python tools/adjoint_bench.py d3q19 float 256 256 256 10 2>&1 | tail -1
