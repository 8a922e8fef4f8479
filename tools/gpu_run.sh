#!/bin/bash
# One parameterised helper for GPU-box runs (called through gpurun): tools/gpu_run.sh <job> [args...]
#   tests                    pytest -m gpu
#   bench [bench.py args]    headline bench line into gpurun_out/
#   ab_kernel                A/B of the stage kernels (classic / persist / ring) on the headline case
set -u
mkdir -p gpurun_out
job=${1:-tests}; shift || true
case "$job" in
tests)
    python -m pytest tests -m gpu -x -q "$@" 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log ;;
bench)
    python bench.py "$@" 2> gpurun_out/bench.err | tee gpurun_out/bench.json ;;
ab_kernel)
    : > gpurun_out/ab_kernel.txt
    for g in "" "--no-globals"; do
      for k in classic persist ring; do
        echo "== TCLB_KERNEL=$k $g" | tee -a gpurun_out/ab_kernel.txt
        TCLB_KERNEL=$k python bench.py --steps 30 --warmup 5 $g "$@" 2>> gpurun_out/ab_kernel.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'], d['clocks'])" | tee -a gpurun_out/ab_kernel.txt
      done
    done ;;
*)
    echo "unknown job $job"; exit 2 ;;
esac
