#!/bin/bash
# strong-scaling pass on one box: N = 1,2,4,8 (up to $1), fused vs NCCL baseline, plus the other named configs
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L | head -8
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["n_gpus"], d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"], d["config"]["parallelism"], d["config"]["halo"])'
echo "== multi-gpu pytest"; timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "multi_gpu or virtual_ranks" 2>&1 | tail -3
launcher() { if [ $1 -eq 1 ]; then echo "python"; else echo "python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2"; fi; }
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    echo "== headline N=$n fused"
    timeout 900 $(launcher $n 2951$n) bench.py --gpus $n --steps 30 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "rc=$?"; cat gpurun_out/scale_n$n.json | python -c "$P"; grep -v "OMP_NUM\|\*\*\*\*\|^$" gpurun_out/scale_n$n.err | tail -3
  fi
done
if [ $N -ge 8 ]; then
  echo "== headline N=8 nccl baseline"
  timeout 900 $(launcher 8 29530) bench.py --gpus 8 --steps 10 --warmup 3 --halo nccl > gpurun_out/scale_nccl_n8.json 2> gpurun_out/scale_nccl_n8.err; echo "rc=$?"; cat gpurun_out/scale_nccl_n8.json | python -c "$P"
  echo "== headline N=8 fused, no Globals"
  timeout 900 $(launcher 8 29531) bench.py --gpus 8 --steps 30 --warmup 5 --no-globals 2>/dev/null | tee gpurun_out/scale_noglob_n8.json | python -c "$P"
fi
echo "== d3q19 512^3 strong scaling (BASELINE config 2)"
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    timeout 600 $(launcher $n 2952$n) bench.py --gpus $n --steps 30 --warmup 5 --model d3q19 --shape 512,512,512 --no-globals > gpurun_out/scale_d3q19_n$n.json 2>/dev/null; cat gpurun_out/scale_d3q19_n$n.json | python -c "$P"
  fi
done
echo "== d3q27_pf_velocity 512^3 (BASELINE config 4; library from the reference model files)"
for n in 1 8; do
  if [ $n -le $N ]; then
    timeout 600 $(launcher $n 2953$n) bench.py --gpus $n --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 512,512,512 --no-globals > gpurun_out/scale_pf_n$n.json 2>/dev/null; cat gpurun_out/scale_pf_n$n.json | python -c "$P"
  fi
done
echo "== weak scaling d3q27_cumulant 256^3 per GPU (BASELINE config 5)"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --shape 256,256,256 --no-globals 2>/dev/null | tee gpurun_out/weak_n1.json | python -c "$P"
if [ $N -ge 8 ]; then timeout 600 $(launcher 8 29541) bench.py --gpus 8 --steps 20 --warmup 5 --shape 256,1024,512 --no-globals 2>/dev/null | tee gpurun_out/weak_n8.json | python -c "$P"; fi
