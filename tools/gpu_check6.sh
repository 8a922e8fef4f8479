#!/bin/bash
# one-GPU check: tests, headline bench, diagnostics, launch list and ncu captures (plain run first, then ncu)
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"], d["clocks"])'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -4 gpurun_out/pytest_gpu.log
echo "== bench N=1" ; timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"; cat gpurun_out/bench_n1.json | python -c "$P"; tail -3 gpurun_out/bench_n1.err
echo "== no globals"; timeout 600 python bench.py --steps 15 --warmup 3 --no-globals 2>/dev/null | tee gpurun_out/bench_noglob.json | python -c "$P"
echo "== d3q19"; timeout 600 python bench.py --steps 15 --warmup 3 --model d3q19 --shape 512,512,512 2>/dev/null | tee gpurun_out/bench_d3q19.json | python -c "$P"
echo "== d2q9"; timeout 600 python bench.py --steps 15 --warmup 3 --model d2q9 --shape 8192,8192,1 2>/dev/null | tee gpurun_out/bench_d2q9.json | python -c "$P"
echo "== d2q9 noglob"; timeout 600 python bench.py --steps 15 --warmup 3 --model d2q9 --shape 8192,8192,1 --no-globals 2>/dev/null | python -c "$P"
echo "== d3q19 noglob"; timeout 600 python bench.py --steps 15 --warmup 3 --model d3q19 --shape 512,512,512 --no-globals 2>/dev/null | python -c "$P"
echo "== ncu"
CMD="python bench.py --steps 3 --warmup 3 --shape 512,256,256"
$CMD > gpurun_out/plain_a.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/plain_b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 3 -c 1 -o gpurun_out/prof_cumulant9 $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
CMD="python bench.py --steps 3 --warmup 3 --shape 512,256,256 --no-globals"
$CMD > gpurun_out/plain_c.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 3 -c 1 -o gpurun_out/prof_cumulant9_noglob $CMD > gpurun_out/ncu_full2.log 2>&1
tail -2 gpurun_out/ncu_full2.log
CMD="python bench.py --steps 3 --warmup 3 --shape 512,512,256 --model d3q19 --no-globals"
$CMD > gpurun_out/plain_d.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 3 -c 1 -o gpurun_out/prof_d3q19c $CMD > gpurun_out/ncu_full3.log 2>&1
tail -2 gpurun_out/ncu_full3.log
