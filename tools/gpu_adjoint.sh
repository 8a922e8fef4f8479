#!/bin/bash
# adjoint on the GPU: tests, throughput next to the primal, one ncu capture of the adjoint stage kernel
mkdir -p gpurun_out
echo "== pytest adjoint"; timeout 900 python -m pytest tests/test_adjoint.py -m gpu -x -q 2>&1 | tail -3
echo "== adjoint throughput"
timeout 600 python tools/adjoint_bench.py d2q9 float 2048 2048 1 20 2>&1 | tail -1
timeout 600 python tools/adjoint_bench.py d2q9 double 2048 2048 1 20 2>&1 | tail -1
timeout 600 python tools/adjoint_bench.py d3q19 float 256 256 256 10 2>&1 | tail -1
CMD="python tools/adjoint_bench.py d2q9 float 2048 2048 1 4"
$CMD > gpurun_out/adj_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_adjoint_stage -s 1 -c 1 -o gpurun_out/prof_adjoint_d2q9 $CMD > gpurun_out/ncu_adj.log 2>&1
tail -2 gpurun_out/ncu_adj.log
