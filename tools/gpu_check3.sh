#!/bin/bash
mkdir -p gpurun_out
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -8 gpurun_out/pytest_gpu.log
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"], d["e2e"]["value"])'
echo "== bench" ; timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_n1.json | python -c "$P"; tail -5 gpurun_out/bench_n1.err
echo "== no globals"; timeout 600 python bench.py --steps 15 --warmup 3 --no-globals 2>/dev/null | python -c "$P"
for m in d3q27_cumulant_mb6 d3q27_cumulant_mb8; do
echo "== $m"; timeout 600 python bench.py --steps 15 --warmup 3 --model $m 2>/dev/null | python -c "$P"
echo "== $m no globals"; timeout 600 python bench.py --steps 15 --warmup 3 --model $m --no-globals 2>/dev/null | python -c "$P"
done
echo "== d3q19"; timeout 600 python bench.py --steps 15 --warmup 3 --model d3q19 --shape 512,512,512 2>/dev/null | python -c "$P"
echo "== d3q19 noglob"; timeout 600 python bench.py --steps 15 --warmup 3 --model d3q19 --shape 512,512,512 --no-globals 2>/dev/null | python -c "$P"
echo "== d2q9"; timeout 600 python bench.py --steps 15 --warmup 3 --model d2q9 --shape 8192,8192,1 2>/dev/null | python -c "$P"
echo "== d2q9 noglob"; timeout 600 python bench.py --steps 15 --warmup 3 --model d2q9 --shape 8192,8192,1 --no-globals 2>/dev/null | python -c "$P"
echo "== ncu"
CMD="python bench.py --steps 3 --warmup 1 --shape 512,256,256"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 2 -c 2 -o gpurun_out/prof_cumulant3 $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
CMD="python bench.py --steps 3 --warmup 1 --shape 512,256,256 --no-globals"
$CMD > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 2 -c 1 -o gpurun_out/prof_cumulant3_noglob $CMD > gpurun_out/ncu_full2.log 2>&1
tail -2 gpurun_out/ncu_full2.log
