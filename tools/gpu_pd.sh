val() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['fraction_of_measured_hbm'], d['clocks']['reasons'])"; }
for m in $LIBS; do echo "== $m 512^3 no globals"; python bench.py --model $m --shape 512,512,512 --no-globals --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done
for m in $LIBS; do echo "== $m 512^3 globals every iteration"; python bench.py --model $m --shape 512,512,512 --steps 20 --warmup 5 2>> gpurun_out/bench.err | val; done
