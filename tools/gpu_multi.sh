#!/bin/bash
# One multi-GPU box visit (gpurun --gpus N -- bash tools/gpu_multi.sh N): the multi-GPU tests, the headline bench of both
# arms under torchrun, the kernel A/B at N ranks, and the case-file runner (one process per GPU against one process for all).
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi -L | tee gpurun_out/multi_gpus.txt
python -m pytest tests/test_gpu_multiprocess.py tests/test_gpu_kernels.py -m gpu -q -k "two_gpus or multi_gpu" 2>&1 | tail -15 | tee gpurun_out/pytest_multi_n$N.log
timeout 900 $TR --master-port 29701 bench.py --impl reference --gpus $N --steps 20 --warmup 5 2> gpurun_out/ref_n$N.err | tee gpurun_out/ref_n$N.json | cut -c1-500
timeout 600 $TR --master-port 29702 bench.py --gpus $N --steps 30 --warmup 5 2> gpurun_out/ours_n$N.err | tee gpurun_out/ours_n$N.json | cut -c1-500
TCLB_KERNEL=classic timeout 600 $TR --master-port 29703 bench.py --gpus $N --steps 30 --warmup 5 2> gpurun_out/ours_classic_n$N.err | tee gpurun_out/ours_classic_n$N.json | cut -c1-300
timeout 600 $TR --master-port 29704 bench.py --gpus $N --steps 30 --warmup 5 --no-globals 2>> gpurun_out/ours_n$N.err | tee gpurun_out/ours_noglobals_n$N.json | cut -c1-300
timeout 600 $TR --master-port 29705 bench.py --config weak --gpus $N --steps 30 --warmup 5 2>> gpurun_out/ours_n$N.err | tee gpurun_out/weak_n$N.json | cut -c1-300
timeout 600 $TR --master-port 29706 bench.py --config d3q19 --gpus $N --steps 30 --warmup 5 2>> gpurun_out/ours_n$N.err | tee gpurun_out/d3q19_n$N.json | cut -c1-300
timeout 600 $TR --master-port 29707 bench.py --config pf_velocity --gpus $N --steps 20 --warmup 5 2>> gpurun_out/ours_n$N.err | tee gpurun_out/pf_n$N.json | cut -c1-300
# the headline XML case through the case-file runner: one process per GPU, then one process driving all GPUs
python tools/case_throughput.py --gpus $N 2>&1 | tail -6 | tee gpurun_out/case_throughput_n$N.txt
