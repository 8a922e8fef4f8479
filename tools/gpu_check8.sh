#!/bin/bash
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["config"]["lattice"], d["value"], d["ms_per_step"], d["roofline"]["bytes_per_update"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"])'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -4 gpurun_out/pytest_gpu.log
echo "== pf 256^3 noglob"; timeout 900 python bench.py --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 256,256,256 --no-globals 2> gpurun_out/pf.err | python -c "$P"; tail -3 gpurun_out/pf.err
echo "== pf 512^3 noglob"; timeout 900 python bench.py --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 512,512,512 --no-globals 2> gpurun_out/pf.err | tee gpurun_out/bench_pf512.json | python -c "$P"; tail -3 gpurun_out/pf.err
echo "== pf 512^3 globals"; timeout 900 python bench.py --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 512,512,512 2> gpurun_out/pf.err | python -c "$P"; tail -3 gpurun_out/pf.err
echo "== ref d3q27_cumulant (reference model files) noglob"; timeout 900 python bench.py --steps 10 --warmup 3 --model ref_d3q27_cumulant --no-globals 2> gpurun_out/pf.err | python -c "$P"; tail -3 gpurun_out/pf.err
echo "== ref d3q19 noglob"; timeout 900 python bench.py --steps 10 --warmup 3 --model ref_d3q19 --shape 512,512,512 --no-globals 2> gpurun_out/pf.err | python -c "$P"; tail -3 gpurun_out/pf.err
echo "== headline"; timeout 900 python bench.py --steps 30 --warmup 5 2>/dev/null | tee gpurun_out/bench_n1.json | python -c "$P"
echo "== ncu pf"
CMD="python bench.py --steps 2 --warmup 3 --shape 256,256,128 --model ref_d3q27_pf_velocity --no-globals"
$CMD > gpurun_out/plain_pf.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_pf.csv $CMD > gpurun_out/ncu_list_pf.log 2>&1
$CMD > gpurun_out/plain_pf2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 12 -c 4 -o gpurun_out/prof_pf $CMD > gpurun_out/ncu_full_pf.log 2>&1
tail -2 gpurun_out/ncu_full_pf.log
