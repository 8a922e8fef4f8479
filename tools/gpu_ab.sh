#!/bin/bash
# A/B of kernel variants (and of an older tree built under build/oldtree) on the same GPU
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"])'
for rep in 1 2; do
for m in d3q27_cumulant_v1 d3q27_cumulant_nodyn; do
echo "== $m noglob"; python bench.py --steps 30 --warmup 10 --no-globals --no-graph --model $m 2>/dev/null | python -c "$P"
done
echo "== OLD noglob"; (cd build/oldtree && python bench.py --steps 30 --warmup 10 --no-globals 2>/dev/null | python -c "$P")
done
for m in d3q27_cumulant_v1 d3q27_cumulant_nodyn; do
echo "== $m globals"; python bench.py --steps 30 --warmup 10 --model $m 2>/dev/null | python -c "$P"
done
echo "== OLD globals"; (cd build/oldtree && python bench.py --steps 30 --warmup 10 2>/dev/null | python -c "$P")
