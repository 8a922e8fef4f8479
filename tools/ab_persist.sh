#!/bin/bash
# A/B of the stage kernels on the headline case (1 GPU): one block per tile (classic) against the persistent kernel with
# the bulk-copy ring at several occupancies (libraries built by tools/build_variants.py ... pbN -DTCLB_PERSIST_MIN_BLOCKS=N)
mkdir -p gpurun_out
: > gpurun_out/ab_kernel2.txt
run() { echo "== $*" | tee -a gpurun_out/ab_kernel2.txt; env $1 python bench.py --steps 30 --warmup 5 ${@:2} 2>> gpurun_out/ab_kernel2.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'], d['clocks']['reasons'])" | tee -a gpurun_out/ab_kernel2.txt; }
for g in "" "--no-globals"; do
  run TCLB_KERNEL=classic $g
  for lib in ${LIBS:-d3q27_cumulant_pb5 d3q27_cumulant_pb6 d3q27_cumulant_pb7}; do
    run TCLB_KERNEL=ring --model $lib $g
  done
done
