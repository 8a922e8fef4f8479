#!/bin/bash
# gpu_final.sh plus the models lowered from the reference's own files (after the C() block rewriting)
bash tools/gpu_final.sh
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["config"]["lattice"], d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"])'
for m in ref_d3q19 ref_d3q27_cumulant ref_d2q9; do
  sh=512,512,512; [ $m = ref_d2q9 ] && sh=8192,8192,1
  echo "== $m"; timeout 600 python bench.py --steps 20 --warmup 5 --model $m --shape $sh --no-globals 2>gpurun_out/$m.err | python -c "$P" || tail -3 gpurun_out/$m.err
done
echo "== ref_d3q27_pf_velocity"; timeout 600 python bench.py --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 512,512,512 --no-globals 2>gpurun_out/pf.err | python -c "$P" || tail -3 gpurun_out/pf.err
