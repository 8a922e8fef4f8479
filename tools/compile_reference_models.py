#!/usr/bin/env python
"""Lower EVERY model directory of a TCLB tree and check that the emitted source compiles (host pass of g++, syntax only;
the same node code is what nvcc sees for the device).  Usage: compile_reference_models.py [root] [model ...]"""
import concurrent.futures as cf
import os
import subprocess
import sys
import traceback

sys.path.insert(0, ".")
from tclb_b200 import build  # noqa: E402
from tclb_b200.gen import rfrontend  # noqa: E402
from tclb_b200.gen.emit import emit_model_source  # noqa: E402

AD = "--ad" in sys.argv          # also check the adjoint (dual-number) build of every model that compiles
args = [a for a in sys.argv[1:] if a != "--ad"]
root = args[0] if args else "/root/reference"
only = args[1:]
out = build.ROOT / "build" / "gen_try"
out.mkdir(parents=True, exist_ok=True)
OPTS = {"d2q9_scmp": {"Kupershtokh": True}, "d2q9_reaction_diffusion_system": {"AllenCahn": True}, "d3q27_PSM": {"NEBB": True}}
jobs = []
status = {}
for d, _, files in sorted(os.walk(os.path.join(root, "models"))):
    if "conf.mk" not in files:
        continue
    name = os.path.basename(d)
    if only and name not in only:
        continue
    try:
        m = rfrontend.load_model(d, options=OPTS.get(name))
        src = out / f"try_{name}.cu"
        src.write_text(emit_model_source(m, "try_" + name, adjoint=AD))
        jobs.append((name, src))
    except Exception as e:
        status[name] = f"lowering FAILED: {type(e).__name__}: {str(e)[:140]}"
        if only:
            traceback.print_exc()


def compile_one(job):
    name, src = job
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-fopenmp", "-x", "c++", "-w", "-include", "cmath", "-include", "math.h", "-include", "cassert", "-DTCLB_DOUBLE", "-I", str(build.ROOT / "include"),
           "-I", str(build.ROOT / "csrc" / "kernels"), "-I", "/usr/local/cuda/include", str(src)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode == 0:
        if AD:
            r2 = subprocess.run(cmd[:-1] + ["-DTCLB_AD", cmd[-1]], capture_output=True, text=True)
            if r2.returncode != 0:
                errs = [l for l in r2.stderr.splitlines() if "error" in l]
                return name, "ok (adjoint build FAILED: " + (errs[0][-140:] if errs else "?") + ")"
            return name, "ok + adjoint build"
        return name, "ok"
    errs = [l for l in r.stderr.splitlines() if "error" in l]
    return name, f"compile FAILED ({len(errs)} errors): " + (errs[0][-170:] if errs else r.stderr[-170:])


with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
    for name, st in ex.map(compile_one, jobs):
        status[name] = st
good = 0
for name in sorted(status):
    print(f"{name:40s} {status[name]}")
    good += status[name].startswith("ok")
print(good, "of", len(status), "models lower and compile")
if AD:
    print(sum(1 for v in status.values() if v == "ok + adjoint build"), "of them also compile as an adjoint (dual-number) build")
