#!/bin/bash
N=${1:-1}
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["n_gpus"], d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"], d["config"]["parallelism"], d["config"]["halo"])'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -4 gpurun_out/pytest_gpu.log
echo "== bench N=1" ; timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"; cat gpurun_out/bench_n1.json | python -c "$P"; tail -3 gpurun_out/bench_n1.err
for m in d3q27_cumulant_mb6 d3q27_cumulant_mb7; do
echo "== $m"; timeout 600 python bench.py --steps 15 --warmup 3 --model $m 2>/dev/null | python -c "$P"
done
echo "== no globals"; timeout 600 python bench.py --steps 15 --warmup 3 --no-globals 2>/dev/null | python -c "$P"
echo "== d3q19"; timeout 600 python bench.py --steps 15 --warmup 3 --model d3q19 --shape 512,512,512 2>/dev/null | python -c "$P"
echo "== d3q19 noglob"; timeout 600 python bench.py --steps 15 --warmup 3 --model d3q19 --shape 512,512,512 --no-globals 2>/dev/null | python -c "$P"
echo "== d2q9"; timeout 600 python bench.py --steps 15 --warmup 3 --model d2q9 --shape 8192,8192,1 2>/dev/null | python -c "$P"
for n in 2 4 8; do
  if [ $n -le $N ]; then
    echo "== bench N=$n fused"
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $n --steps 30 --warmup 5 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "rc=$?"; cat gpurun_out/bench_n$n.json | python -c "$P"; tail -2 gpurun_out/bench_n$n.err
    echo "== bench N=$n nccl baseline"
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $n --steps 10 --warmup 3 --halo nccl > gpurun_out/bench_nccl_n$n.json 2> gpurun_out/bench_nccl_n$n.err; echo "rc=$?"; cat gpurun_out/bench_nccl_n$n.json | python -c "$P"; tail -2 gpurun_out/bench_nccl_n$n.err
  fi
done
echo "== ncu"
CMD="python bench.py --steps 3 --warmup 1 --shape 512,256,256"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 2 -c 1 -o gpurun_out/prof_cumulant4 $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
CMD="python bench.py --steps 3 --warmup 1 --shape 512,512,256 --model d3q19 --no-globals"
$CMD > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 2 -c 1 -o gpurun_out/prof_d3q19 $CMD > gpurun_out/ncu_full2.log 2>&1
tail -2 gpurun_out/ncu_full2.log
