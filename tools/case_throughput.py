#!/usr/bin/env python
"""Throughput of the headline XML case through the case-file runner on N GPUs:
    (a) one process driving all N GPUs   bin/tclb --gpu --ranks N d3q27_cumulant case.xml
    (b) one process per GPU              python -m torch.distributed.run --nproc-per-node N -m tclb_b200.run --gpu ...
Both write the reference-format Log CSV; the rate is taken from its Walltime column between iteration W and W+K
(Log every iteration = Globals every iteration, as in bench.py)."""
import argparse
import csv
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASE = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/">
  <Geometry nx="{nx}" ny="{ny}" nz="{nz}">
    <MRT><Box/></MRT>
    <Wall mask="ALL"><Box ny="1"/><Box dy="-1"/></Wall>
  </Geometry>
  <Model>
    <Param name="nu" value="0.005"/>
    <Param name="ForceX" value="1e-6"/>
    <Param name="Velocity" value="0.05"/>
  </Model>
  <Log Iterations="1"/>
  <Solve Iterations="{iters}"/>
</CLBConfig>
"""


def rate(logfile, w, k, nodes):
    rows = list(csv.reader(open(logfile)))
    it = rows[0].index("Iteration")
    wt = rows[0].index("Walltime")
    t = {int(float(r[it])): float(r[wt]) for r in rows[1:]}
    return nodes * k / (t[w + k] - t[w]) / 1e6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--shape", default="1024,512,512")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    a = ap.parse_args()
    nx, ny, nz = (int(v) for v in a.shape.split(","))
    nodes = float(nx) * ny * nz
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    for label, cmd in (
        ("one process, %d GPUs" % a.gpus, [os.path.join(ROOT, "bin", "tclb"), "--gpu", "--precision", "float", "--ranks", str(a.gpus), "d3q27_cumulant", "case.xml"]),
        ("%d processes, one GPU each" % a.gpus, [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus), "--master-addr", "127.0.0.1",
                                                  "--master-port", "29741", "-m", "tclb_b200.run", "--gpu", "--precision", "float", "d3q27_cumulant", "case.xml"]),
    ):
        wd = tempfile.mkdtemp(prefix="tclb_case_")
        with open(os.path.join(wd, "case.xml"), "w") as f:
            f.write(CASE.format(nx=nx, ny=ny, nz=nz, iters=a.steps + a.warmup + 1))
        r = subprocess.run(cmd, cwd=wd, env=env, capture_output=True, text=True, timeout=900)
        if r.returncode != 0:
            print(label, "FAILED", (r.stdout + r.stderr)[-1500:])
            continue
        logs = [f for f in os.listdir(os.path.join(wd, "output")) if "_Log_" in f]
        print("%-32s %10.1f MLUPS  (Log CSV Walltime, iterations %d..%d)" % (label, rate(os.path.join(wd, "output", logs[0]), a.warmup, a.steps, nodes), a.warmup, a.warmup + a.steps))


if __name__ == "__main__":
    main()
