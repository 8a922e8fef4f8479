#!/bin/bash
# end-of-round verification on one GPU: tests, smoke, headline bench, diagnostics
mkdir -p gpurun_out
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["model"], d["config"]["lattice"], d["value"], d["ms_per_step"], d["roofline"]["fraction_of_measured_hbm"], "e2e", d["e2e"]["value"], "graph", d.get("graph_iterations"), d["clocks"]["reasons"])'
echo "== pytest" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -4 gpurun_out/pytest_gpu.log
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== headline (default flags)"; timeout 900 python bench.py > gpurun_out/bench_default.json 2>gpurun_out/h.err; echo "rc=$?"; python -c "$P" < gpurun_out/bench_default.json; tail -2 gpurun_out/h.err
echo "== reference arm"; python bench.py --impl reference
echo "== noglob graph"; timeout 900 python bench.py --steps 30 --warmup 10 --no-globals 2>/dev/null | python -c "$P"
echo "== noglob plain"; timeout 900 python bench.py --steps 30 --warmup 10 --no-globals --no-graph 2>/dev/null | python -c "$P"
echo "== d3q19"; timeout 900 python bench.py --steps 30 --warmup 10 --model d3q19 --shape 512,512,512 --no-globals 2>/dev/null | python -c "$P"
echo "== profile table"; cat > /tmp/k.xml <<'XML'
<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/">
  <Geometry nx="256" ny="128" nz="128"><MRT><Box/></MRT><Wall mask="ALL"><Box ny="1"/><Box dy="-1"/></Wall></Geometry>
  <Model><Param name="nu" value="0.01"/><Param name="ForceX" value="1e-6"/></Model>
  <Log Iterations="50"/>
  <Solve Iterations="100"/>
</CLBConfig>
XML
(cd /tmp && TCLB_PROFILE=1 $GRAFT_REPO_ROOT/bin/tclb --gpu --precision float d3q27_cumulant k.xml 2>&1 | grep -A6 "launches" | head -8)
