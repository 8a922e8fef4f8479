mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/pytest_gpu.log; tail -n 5 gpurun_out/pytest_gpu.log
python bench.py --steps 30 --warmup 5 2> gpurun_out/bench.err | tee gpurun_out/bench_headline.json | cut -c1-160
LIBS="ref_d3q27_pf_velocity ref_d3q27_pf_velocity_mb4 ref_d3q27_pf_velocity_mb5" bash tools/gpu_pd.sh
val() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['fraction_of_measured_hbm'], d['clocks']['reasons'])"; }
for m in ref_d3q27_cumulant ref_d3q19 d3q19 d3q27_cumulant; do echo "== $m 512^3 no globals (classic kernel)"; python bench.py --model $m --shape 512,512,512 --no-globals --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done | tee gpurun_out/ref_lowered.txt
python bench.py --config d3q19 --steps 30 --warmup 5 2>> gpurun_out/bench.err | tee gpurun_out/bench_d3q19.json | cut -c1-120
