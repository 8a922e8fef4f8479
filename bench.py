#!/usr/bin/env python
"""Headline benchmark: whole-job MLUPS of the d3q27_cumulant model with Globals reduced every iteration.

Config (BASELINE.json configs[2] + north star): d3q27_cumulant, fp32, turbulent-channel geometry
1024x512x512 (walls on the y faces, body force, synthetic data), force/flux Globals every iteration,
STRONG scaling over N GPUs of one node (y/z decomposition, the reference's rule).  The metric is the
reference's own (src/main.cpp:101-127): MLBUps = global padded nx*ny*nz * iterations / time.

  python bench.py --gpus 1 --steps 50 --warmup 5
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P bench.py --gpus 8 ...
  python bench.py --impl reference          # the unmodified reference arm (not installable here)

Timed region: exactly K iterations, CUDA events on the launching stream, barrier + synchronize on
both sides, MAX over ranks.  The two 29 GB snapshots are far larger than the 126 MB L2, so no flush
is needed between iterations (stated in `config`).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODEL = "d3q27_cumulant"
SHAPE = (1024, 512, 512)

# BASELINE.json configs[1..4]; `--config cumulant` (configs[2]) is the headline and the default.
#   walls: which faces are Wall nodes; globals: reduce the Globals every iteration (else never inside the timed loop)
CONFIGS = {
    "cumulant": dict(model="d3q27_cumulant", shape=(1024, 512, 512), walls="y", globals=True, scaling="strong",
                     case="turbulent channel (y walls, body force), Globals every iteration"),
    "d3q19": dict(model="d3q19", shape=(512, 512, 512), walls="xyz", globals=False, scaling="strong",
                  case="closed Wall box driven by a body force (stand-in for the lid-driven cavity: the reference's d3q19 has no "
                       "moving-wall node type, SURVEY 7.4)"),
    "pf_velocity": dict(model="ref_d3q27_pf_velocity", shape=(512, 512, 512), walls="y", globals=False, scaling="strong",
                        case="phase-field model lowered from the reference's own model files, 4-stage iteration, two lattices"),
    "weak": dict(model="d3q27_cumulant", per_gpu=256, walls="y", globals=True, scaling="weak",
                 case="weak-scaling halo sweep: PER_GPU^3 nodes per GPU (--per-gpu 128..384), Globals every iteration"),
}


def weak_shape(per_gpu: int, n: int):
    """global lattice that gives every rank a per_gpu^3 block under the reference's decomposition rule
    (y*z split minimising divz*ny + divy*nz, src/Solver.cpp.Rt:288-350): 2 -> 2x1, 4 -> 2x2, 8 -> 4x2 (y x z)"""
    dy, dz = {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}.get(n, (n, 1))
    return (per_gpu, per_gpu * dy, per_gpu * dz)


REF_CASE = """<?xml version="1.0"?>
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
  {log}
  <Solve Iterations="{iters}"/>
</CLBConfig>
"""


def reference_arm(a):
    """The UNMODIFIED reference (nkkko/TCLB) built by baseline/build_reference.py: its own kernels, engine, handlers and
    XML front-end; the only substitutes are the generator (this repository's R interpreter runs the reference's
    src/conf.R and *.Rt templates, there is no R here) and the MPI library (baseline/mpi_stub, single node).
    Same case as the main arm: d3q27_cumulant 1024x512x512, walls on the y faces, body force, <Log Iterations="1"/> =
    Globals reduced and logged every iteration (src/Handlers/cbLog.cpp:14-15).  Timed with the reference's own clock:
    the Walltime column of its Log CSV between iteration W and W+K (host time including its per-iteration
    synchronisation, Globals read-back and log write: that IS the reference's hot path in this configuration)."""
    import csv
    import shutil
    import tempfile
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    exe = os.path.join(ROOT, "baseline", "_ref", "CLB", MODEL, "main")

    def unavailable(why):
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": why}))
        return 0
    if a.model != MODEL:
        return unavailable(f"the reference arm is built for {MODEL} only")
    if not os.path.exists(exe):
        if rank == 0:
            r = subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "build_reference.py"), MODEL], capture_output=True, text=True)
            if r.returncode != 0 or not os.path.exists(exe):
                return unavailable("baseline/build_reference.py failed here: " + (r.stdout + r.stderr).strip().splitlines()[-1][:300])
        else:
            for _ in range(1200):
                if os.path.exists(exe):
                    break
                time.sleep(0.5)
    shape = tuple(int(v) for v in a.shape.split(",")) if a.shape else SHAPE
    nx, ny, nz = shape
    tag = f"{os.environ.get('MASTER_PORT', '0')}_{os.getppid() if world > 1 else os.getpid()}"
    wd = os.path.join(tempfile.gettempdir(), f"tclb_ref_{tag}")
    if rank == 0:
        shutil.rmtree(wd, ignore_errors=True)
        os.makedirs(os.path.join(wd, "output"), exist_ok=True)
        log = "" if a.no_globals else '<Log Iterations="1"/>'
        iters = a.warmup + a.steps
        with open(os.path.join(wd, "bench.xml.tmp"), "w") as f:
            f.write(REF_CASE.format(nx=nx, ny=ny, nz=nz, iters=iters, log=log))
        os.replace(os.path.join(wd, "bench.xml.tmp"), os.path.join(wd, "bench.xml"))
    else:
        for _ in range(600):
            if os.path.exists(os.path.join(wd, "bench.xml")):
                break
            time.sleep(0.1)
    env = dict(os.environ, TCLB_MPI_RANK=str(rank), TCLB_MPI_SIZE=str(world), TCLB_MPI_ID=tag)
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    sampler.start()
    t0 = time.perf_counter()
    r = subprocess.run([exe, "bench.xml"], cwd=wd, env=env, capture_output=True, text=True)
    wall = time.perf_counter() - t0
    sampler.stop_evt.set()
    sampler.join()
    if r.returncode != 0:
        sys.stderr.write(r.stdout[-3000:] + r.stderr[-3000:])
        return unavailable(f"reference main exited with {r.returncode}: " + (r.stdout + r.stderr).strip().splitlines()[-1][:300])
    if rank != 0:
        return 0
    nodes = float(nx) * ny * nz
    out = {"metric": "MLUPS (whole box, device-timed, max over ranks)", "unit": "MLUPS", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "reference",
           "config": {"model": MODEL, "case": "turbulent channel (y walls, body force), Log every iteration" if not a.no_globals else "no globals",
                      "lattice": list(shape), "global_batch": int(nodes), "seq_len": None, "parallelism": f"reference MPIDivision over {world} rank(s)",
                      "mpi": "baseline/mpi_stub (shared-memory stand-in, single node)", "generator": "reference src/*.Rt expanded by tclb_b200.gen (no R in the image)",
                      "l2": "working set 2x%.1f GB >> 126 MB L2" % (nodes * 27 * 4 / 1e9)},
           "clocks": sampler.result(), "reference_wall_s": round(wall, 2)}
    logs = [f for f in os.listdir(os.path.join(wd, "output")) if "_Log_" in f and f.endswith(".csv")]
    if logs and not a.no_globals:
        rows = list(csv.reader(open(os.path.join(wd, "output", logs[0]))))
        hdr = [h.strip().strip('"') for h in rows[0]]
        it, wt = hdr.index("Iteration"), hdr.index("Walltime")
        t = {int(float(x[it])): float(x[wt]) for x in rows[1:] if len(x) > wt}
        w0, w1 = a.warmup, a.warmup + a.steps
        if w0 in t and w1 in t and t[w1] > t[w0]:
            sec = t[w1] - t[w0]
            out.update(value=round(nodes * a.steps / sec / 1e6, 1), ms_per_step=round(sec * 1e3 / a.steps, 4), timing="reference Log CSV Walltime column, iterations W..W+K")
            out["e2e"] = {"value": out["value"], "unit": "MLUPS", "note": "the reference's per-iteration path already includes its Globals read-back and log write"}
    if "value" not in out:
        # no per-iteration log: the reference's own monitor line (src/main.cpp:101-127) over the whole Solve segment
        import re
        mm = re.findall(r"([0-9.]+) MLBUps", r.stdout)
        if not mm:
            return unavailable("reference ran but printed no MLBUps line")
        out.update(value=float(mm[-1]), ms_per_step=round(nodes / float(mm[-1]) / 1e3, 4), timing="reference monitor line (CUDA events)")
    out["gpu_launches"] = 2 * (a.warmup + a.steps)
    print(json.dumps(out))
    shutil.rmtree(wd, ignore_errors=True)
    return 0


class ClockSampler(threading.Thread):
    """nvidia-smi clocks/throttle reasons sampled during the timed region."""

    def __init__(self, device_index: int):
        super().__init__(daemon=True)
        self.dev = device_index
        self.stop_evt = threading.Event()
        self.samples = []
        self.max_mhz = None
        self.reasons = set()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.dev}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split("\n")[0]
                f = [x.strip() for x in out.split(",")]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            self.stop_evt.wait(0.2)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", default="cumulant", choices=sorted(CONFIGS), help="named benchmark configuration of BASELINE.json")
    ap.add_argument("--per-gpu", type=int, default=None, help="--config weak: edge of the cubic block each GPU owns (128..384)")
    ap.add_argument("--shape", default=None, help="nx,ny,nz override (diagnostics only; the headline uses 1024,512,512)")
    ap.add_argument("--model", default=None)
    ap.add_argument("--no-globals", action="store_true", help="diagnostics: skip the per-iteration Globals")
    ap.add_argument("--halo", default="direct", choices=["direct", "nccl"])
    ap.add_argument("--block", default=None, help="bx,by kernel block override")
    ap.add_argument("--no-graph", action="store_true", help="diagnostics: plain launches instead of the replayed CUDA graph")
    ap.add_argument("--precision", default="float", help="diagnostics: arithmetic[_storage] of the library (headline: float)")
    a = ap.parse_args()
    cfg = CONFIGS[a.config]
    if a.model is None:
        a.model = cfg["model"]
    if a.shape is None and a.config != "cumulant":
        sh = weak_shape(a.per_gpu or cfg["per_gpu"], a.gpus) if cfg["scaling"] == "weak" else cfg["shape"]
        a.shape = ",".join(str(v) for v in sh)
    if not cfg["globals"]:
        a.no_globals = True
    if a.impl == "reference":
        return reference_arm(a)

    import numpy as np
    import torch

    from tclb_b200.parallel.dist import DistDomain
    from tclb_b200.runtime import Model

    # stdout must carry exactly one JSON line: libraries (NCCL prints its version banner there) write to fd 1,
    # so fd 1 is pointed at stderr until the result is ready
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world}"
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the CPU oracle is not a benchmark target)")
    torch.cuda.set_device(local_rank)
    from tclb_b200.runtime import core
    core().tclb_log_init(rank, 6)          # warnings and errors only: stdout carries exactly one JSON line
    shape = tuple(int(v) for v in a.shape.split(",")) if a.shape else SHAPE
    nx, ny, nz = shape
    model = Model(a.model, a.precision)
    dd = DistDomain(model, shape, rank=rank, world=world, device=local_rank, direct=(a.halo == "direct"))
    d = dd.domain
    if a.no_graph:
        d.use_graphs(False)
    if a.block:
        bx, by = (int(v) for v in a.block.split(","))
        d.set_block(bx, by)
    # synthetic turbulent-channel geometry: walls on both y faces, collision everywhere else, body force along x
    x0, y0, z0, bx_, by_, bz_ = d.boxes[rank]
    fl = np.full((bz_, by_, bx_), model.flag("MRT"), dtype=np.uint32)
    if y0 == 0:
        fl[:, 0, :] = model.flag("Wall")
    if y0 + by_ == ny:
        fl[:, -1, :] = model.flag("Wall")
    if "x" in cfg["walls"]:
        fl[:, :, 0] = model.flag("Wall")
        fl[:, :, -1] = model.flag("Wall")
    if "z" in cfg["walls"]:
        if z0 == 0:
            fl[0, :, :] = model.flag("Wall")
        if z0 + bz_ == nz:
            fl[-1, :, :] = model.flag("Wall")
    dd.set_local_flags(fl)
    known = set(model.settings) | set(model.zonesettings)
    zonal_input = "Velocity" if "Velocity" in known else ("VelocityX" if "VelocityX" in known else None)
    for name, val in (("nu", 0.005), ("Viscosity", 0.005), ("ForceX", 1e-6), ("GravitationX", 1e-6), (zonal_input, 0.05)):
        if name in known:
            d.set_setting(name, val)
    d.init()
    use_globals = not a.no_globals
    stream = torch.cuda.ExternalStream(d.stream(rank), device=local_rank)

    def run(n):
        d.iterate(n, globals_on_last=False, globals_every=use_globals)

    dd.barrier()
    run(a.warmup)
    d.sync()
    dd.barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dd.barrier()
    import ctypes
    _mlib = ctypes.CDLL(model.path)          # the model library the core already loaded: its own count of kernel launches
    _mlib.tclb_model_launch_count.restype = ctypes.c_longlong
    launches_before = int(_mlib.tclb_model_launch_count())
    e0.record(stream)
    run(a.steps)
    e1.record(stream)
    launches_timed = int(_mlib.tclb_model_launch_count()) - launches_before
    d.sync()
    torch.cuda.synchronize()
    dd.barrier()
    ms = e0.elapsed_time(e1)
    sampler.stop_evt.set()
    sampler.join()
    ms_max = dd.allreduce_max(ms)
    g = d.globals() if use_globals else {}

    # ---- end to end through the public API: every step uploads that step's inputs (the zonal inlet/forcing table,
    # staged in pinned host memory) and reads that step's result (the reduced Globals) back to the host ----
    e2e_steps = max(5, a.steps // 2)
    h2d_bytes, d2h_bytes = dd.e2e_bytes_per_step()
    d.sync()
    dd.barrier()
    t0 = time.perf_counter()
    e0.record(stream)
    last = None
    for i in range(e2e_steps):
        if zonal_input:
            d.set_setting(zonal_input, 0.05 + 1e-4 * (i % 7))  # this step's input -> H2D of the zone table
        d.iterate(1, globals_every=True)                       # fused stream-collide + Globals exchange
        last = d.globals()[model.globals[0]]                   # this step's result -> D2H of the Globals ring slot
    e1.record(stream)
    d.sync()
    torch.cuda.synchronize()
    dd.barrier()
    e2e_ms = dd.allreduce_max((time.perf_counter() - t0) * 1e3)
    nodes = float(nx) * ny * nz
    value = nodes * a.steps / (ms_max * 1e-3) / 1e6
    e2e_value = nodes * e2e_steps / (e2e_ms * 1e-3) / 1e6
    bytes_per_update = model.bytes_per_update
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)
    if rank == 0:
        out = {
            "metric": "MLUPS (whole box, device-timed, max over ranks)", "value": round(value, 1), "unit": "MLUPS",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms_max / a.steps, 4),
            "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None,
            "dtype": "fp32" if a.precision == "float" else a.precision, "data": "synthetic",
            "impl": "ours",
            "config": {"name": a.config, "model": a.model, "case": cfg["case"] if (use_globals or not cfg["globals"]) else "no globals (diagnostic)",
                       "lattice": list(shape), "global_batch": int(nodes), "seq_len": None,
                       "parallelism": f"domain-decomposition y{d.divy}xz{d.divz}", "halo": "fused P2P stores" if a.halo == "direct" else "nccl baseline",
                       "l2": "working set 2x%.1f GB per job >> 126 MB L2, no flush needed" % (nodes * len(model.fields) * model.d.storage_size / 1e9)},
            "clocks": sampler.result(),
            "e2e": {"value": round(e2e_value, 1), "unit": "MLUPS", "steps": e2e_steps, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes},
            "gpu_launches": launches_timed,
            "graph_iterations": d.graph_iterations,
            "roofline": {"bytes_per_update": bytes_per_update, "achieved_gbs_per_gpu": round(value * 1e6 * bytes_per_update / 1e9 / a.gpus, 1),
                         "hbm_gbs_measured": hbm, "fraction_of_measured_hbm": round(value * 1e6 * bytes_per_update / 1e9 / a.gpus / hbm, 3)},
            "globals_sample": {k: g[k] for k in list(g)[:4]}, "last_e2e_flux": last,
        }
        result_line = json.dumps(out)
    else:
        result_line = None
    dd.close()
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    if result_line is not None:
        os.write(1, (result_line + "\n").encode())
    return 0


if __name__ == "__main__":
    sys.exit(main())
