val() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['fraction_of_measured_hbm'], d['clocks']['reasons'])"; }
for m in $LIBS; do echo "== $m pf_velocity 512^3"; python bench.py --config pf_velocity --model $m --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done
