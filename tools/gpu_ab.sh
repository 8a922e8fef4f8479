#!/bin/bash
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["config"]["lattice"], d["value"], d["ms_per_step"], d["clocks"])'
nvidia-smi --query-gpu=pci.bus_id,clocks.mem,clocks.max.mem,power.limit --format=csv
for rep in 1 2; do
echo "== NEW noglob"; python bench.py --steps 30 --warmup 10 --no-globals --no-graph 2>/dev/null | python -c "$P"
echo "== OLD noglob"; (cd build/oldtree && python bench.py --steps 30 --warmup 10 --no-globals 2>/dev/null | python -c "$P")
done
echo "== NEW globals"; python bench.py --steps 30 --warmup 10 2>/dev/null | python -c "$P"
echo "== OLD globals"; (cd build/oldtree && python bench.py --steps 30 --warmup 10 2>/dev/null | python -c "$P")
