#!/bin/bash
# First GPU pass: tests, headline bench, block-shape sweep, launch list and one full ncu capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
which R Rscript mpirun mpicxx autoconf h5cc > gpurun_out/toolchain_probe.txt 2>&1
echo "== pytest" ; timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -15 gpurun_out/pytest_gpu.log
echo "== bench" ; timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_n1.json; tail -5 gpurun_out/bench_n1.err
for blk in 128,1 256,1 128,2 64,2 512,1; do
  echo "== block $blk"; timeout 600 python bench.py --steps 15 --warmup 3 --block $blk 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline'])"
done
echo "== no globals"; timeout 600 python bench.py --steps 15 --warmup 3 --no-globals 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline'])"
echo "== d3q19"; timeout 600 python bench.py --steps 15 --warmup 3 --model d3q19 --shape 512,512,512 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline'])"
echo "== ncu"
CMD="python bench.py --steps 3 --warmup 1 --shape 512,256,256"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stage -s 2 -c 2 -o gpurun_out/prof_cumulant $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
