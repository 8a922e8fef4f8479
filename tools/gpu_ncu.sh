#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_window.py > gpurun_out/window_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/prof_window python tools/ncu_window.py > gpurun_out/ncu_window.log 2>&1
tail -n 3 gpurun_out/window_plain.log; tail -n 3 gpurun_out/ncu_window.log
ls -la gpurun_out/*.ncu-rep
