#!/bin/bash
# One parameterised helper for GPU-box runs (called through gpurun):  tools/gpu_run.sh <job> [args...]
#   tests [pytest args]      pytest -m gpu
#   bench [bench.py args]    one bench line into gpurun_out/bench.json
#   visit                    the whole GPU suite, the headline and the other single-GPU configurations
#   ab LIB...                kernel A/B on the headline case: every library (tools/build_variants.py) with the classic and the ring
#                            kernel, with Globals every iteration and without
#   models LIB...            libraries at 512^3: no Globals (classic kernel) and Globals every iteration
#   pf LIB...                libraries on the pf_velocity configuration
#   ncu                      ONE ncu --set full capture of a window of kernels (tools/ncu_window.py), after its plain run exited 0
#   multi N                  N GPUs: multi-GPU tests, both bench arms under torchrun, the other configurations, the case-file runners
set -u
mkdir -p gpurun_out
job=${1:-tests}; shift || true
val() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'], d['clocks']['reasons'])"; }
case "$job" in
tests)
    python -m pytest tests -m gpu -q "$@" 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log ;;
bench)
    python bench.py "$@" 2> gpurun_out/bench.err | tee gpurun_out/bench.json ;;
visit)
    python -m pytest tests -m gpu -q -s 2>&1 | grep -v "arena\]" | tail -30 > gpurun_out/pytest_gpu.log; tail -n 12 gpurun_out/pytest_gpu.log
    python bench.py --steps 30 --warmup 5 2> gpurun_out/bench.err | tee gpurun_out/bench_headline.json | cut -c1-160
    python bench.py --config d3q19 --steps 30 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_d3q19.json | cut -c1-120
    python bench.py --config pf_velocity --steps 20 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_pf.json | cut -c1-120 ;;
ab)
    : > gpurun_out/ab_kernel2.txt
    for lib in "$@"; do for k in classic ring; do for g in "" "--no-globals"; do
        echo "== TCLB_KERNEL=$k --model $lib $g" | tee -a gpurun_out/ab_kernel2.txt
        TCLB_KERNEL=$k python bench.py --steps 30 --warmup 5 --model $lib $g 2>> gpurun_out/ab_kernel2.err | val | tee -a gpurun_out/ab_kernel2.txt
    done; done; done ;;
models)
    for m in "$@"; do echo "== $m 512^3 no globals"; python bench.py --model $m --shape 512,512,512 --no-globals --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done | tee gpurun_out/models.txt
    for m in "$@"; do echo "== $m 512^3 globals every iteration"; python bench.py --model $m --shape 512,512,512 --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done | tee -a gpurun_out/models.txt ;;
pf)
    for m in "$@"; do echo "== $m pf_velocity 512^3"; python bench.py --config pf_velocity --model $m --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done | tee gpurun_out/pf.txt ;;
ncu)
    python tools/ncu_window.py > gpurun_out/window_plain.log 2>&1 && \
    ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/prof_window python tools/ncu_window.py > gpurun_out/ncu_window.log 2>&1
    tail -n 3 gpurun_out/window_plain.log; tail -n 3 gpurun_out/ncu_window.log; ls -la gpurun_out/*.ncu-rep ;;
multi)
    N=${1:-2}
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
    python tools/case_throughput.py --gpus $N 2>&1 | tail -6 | tee gpurun_out/case_throughput_n$N.txt ;;
*)
    echo "unknown job $job"; exit 2 ;;
esac
