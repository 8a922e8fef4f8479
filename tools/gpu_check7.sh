#!/bin/bash
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"])'
for m in ${MODELS:-d3q27_cumulant d3q27_cumulant_g7 d3q27_cumulant_g6 d3q27_cumulant_pw d3q27_cumulant_pwg6}; do
echo "== $m"; timeout 600 python bench.py --steps 20 --warmup 3 --model $m 2>/dev/null | python -c "$P"
done
