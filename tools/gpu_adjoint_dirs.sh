#!/bin/bash
# chunk width (tangent directions per pass) of the adjoint kernel: throughput for several widths
for d in 3 5 9; do echo "== d2q9 float dirs=$d"; AD_DIRS=$d timeout 900 python tools/adjoint_bench.py d2q9 float 2048 2048 1 20 2>&1 | tail -1; grep -A3 "Compiling entry function.*k_adjoint_stageILi0" build/logs/d2q9_float_ad.ptxas.log | grep -o "Used [0-9]* registers\|[0-9]* bytes spill stores" | tr '\n' ' '; echo; done
for d in 5 7 10; do echo "== d3q19 float dirs=$d"; AD_DIRS=$d timeout 900 python tools/adjoint_bench.py d3q19 float 256 256 256 10 2>&1 | tail -1; grep -A3 "Compiling entry function.*k_adjoint_stageILi0" build/logs/d3q19_float_ad.ptxas.log | grep -o "Used [0-9]* registers\|[0-9]* bytes spill stores" | tr '\n' ' '; echo; done
