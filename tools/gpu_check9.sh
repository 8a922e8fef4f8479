#!/bin/bash
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["config"]["lattice"], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"])'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -4 gpurun_out/pytest_gpu.log
for m in ref_d3q27_pf_velocity ref_d3q27_pf_velocity_mb5 ref_d3q27_pf_velocity_mb6 ref_d3q27_pf_velocity_mb8; do
echo "== $m 512^3 noglob"; timeout 900 python bench.py --steps 10 --warmup 3 --model $m --shape 512,512,512 --no-globals 2> gpurun_out/pf.err | python -c "$P"; tail -3 gpurun_out/pf.err
done
