#!/bin/bash
# 8-GPU confirmation after the front-end change: headline, pf_velocity and ref_d3q19 from the reference's files
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["n_gpus"], d["config"]["lattice"], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["config"]["parallelism"])'
run() { port=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 "$@" 2>gpurun_out/n8_$port.err | tee gpurun_out/n8_$port.json | python -c "$P" || grep -v "OMP_NUM\|\*\*\*\*\|^$" gpurun_out/n8_$port.err | tail -5; }
echo "== headline"; run 29571 --steps 30 --warmup 5
echo "== pf_velocity"; run 29572 --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 512,512,512 --no-globals
echo "== pf_velocity globals"; run 29573 --steps 10 --warmup 3 --model ref_d3q27_pf_velocity --shape 512,512,512
echo "== ref_d3q19"; run 29574 --steps 30 --warmup 5 --model ref_d3q19 --shape 512,512,512 --no-globals
