#!/bin/bash
# strong-scaling pass on one box: N = 1,2,4,8 (up to $1), fused vs NCCL baseline
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L | head -8
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["n_gpus"], d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"], d["config"]["parallelism"], d["config"]["halo"], d["globals_sample"]["Density"])'
echo "== multi-gpu pytest"; timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "multi_gpu or virtual_ranks" 2>&1 | tail -3
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    echo "== bench N=$n fused"
    if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n"; fi
    timeout 900 $L bench.py --gpus $n --steps 30 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "rc=$?"; cat gpurun_out/scale_n$n.json | python -c "$P"; grep -v "OMP_NUM\|\*\*\*\*\|^$" gpurun_out/scale_n$n.err | tail -3
    if [ $n -gt 1 ]; then
    echo "== bench N=$n nccl baseline"
    timeout 900 $L bench.py --gpus $n --steps 10 --warmup 3 --halo nccl > gpurun_out/scale_nccl_n$n.json 2> gpurun_out/scale_nccl_n$n.err; echo "rc=$?"; cat gpurun_out/scale_nccl_n$n.json | python -c "$P"; grep -v "OMP_NUM\|\*\*\*\*\|^$" gpurun_out/scale_nccl_n$n.err | tail -3
    fi
  fi
done
echo "== d3q19 512^3 strong scaling (BASELINE config 2)"
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n"; fi
    timeout 600 $L bench.py --gpus $n --steps 30 --warmup 5 --model d3q19 --shape 512,512,512 --no-globals > gpurun_out/scale_d3q19_n$n.json 2>/dev/null; cat gpurun_out/scale_d3q19_n$n.json | python -c "$P"
  fi
done
