#!/bin/bash
# kernel-variant A/B on the headline case: LIBS="variant ..." (libraries from tools/build_variants.py); every library is run
# with the classic and the ring kernel, with Globals every iteration and without
mkdir -p gpurun_out
: > gpurun_out/ab_kernel2.txt
run() { echo "== $*" | tee -a gpurun_out/ab_kernel2.txt; env $1 python bench.py --steps 30 --warmup 5 ${@:2} 2>> gpurun_out/ab_kernel2.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'], d['clocks']['reasons'])" | tee -a gpurun_out/ab_kernel2.txt; }
for lib in $LIBS; do for k in classic ring; do for g in "" "--no-globals"; do run TCLB_KERNEL=$k --model $lib $g; done; done; done
