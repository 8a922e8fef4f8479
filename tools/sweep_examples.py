#!/usr/bin/env python
"""Run a set of the reference's example cases (example/**.xml) for a few iterations on the CPU backend and report which
ones run: a smoke test of the XML front-end, geometry, handlers and units against real case files.  Edit `cases` to taste.
    python tools/sweep_examples.py
Known and expected failures: examples written for node types their model no longer declares (`Inlet`/`Outlet`/`XYslice1`
in flow/3d, `MRT` in electrokinetic/poison_boltzmann.xml), cases that need their MPMD partner program or files addressed
relative to the reference's root."""
import os, re, subprocess, sys, tempfile, shutil
ROOT="/root/repo"; EX="/root/reference/example"
cases = [
 ("ref_d2q9_npe_guo", "electrokinetic/npe_guo.xml"), ("ref_d2q9_npe_guo", "electrokinetic/npe_guo_nonuniform.xml"), ("ref_d2q9_npe_guo", "electrokinetic/npe_guo_zeta_zero.xml"),
 ("ref_d2q9_poison_boltzmann", "electrokinetic/poison_boltzmann.xml"), ("d2q9_poison_boltzmann", "electrokinetic/poison_boltzmann.xml"),
 ("diffusion2D", "PDE/diffusion2D.xml"), ("wave2D", "PDE/wave2D.xml"), ("advection_diffusion2D", "PDE/advection_diffusion.xml"),
 ("ref_d3q27_pf_velocity", "multiphase/d3q27_pf_velocity/ContactAngle_45.xml"), ("ref_d3q27_pf_velocity", "multiphase/d3q27_pf_velocity/RayleighTaylorInstability_3D.xml"),
 ("ref_d3q27_pf_velocity", "multiphase/d3q27_pf_velocity/annularTaylorBubble_DasC.xml"),
 ("ref_d2q9_plate", "moving/flapping1.xml"),
 ("ref_d3q27_PSM_NEBB", "particle/3d/cube.xml"), ("ref_d3q27_PSM_NEBB", "particle/3d/roll_lb.xml"),
]
import glob
bad = 0
for model, rel in cases:
    src = os.path.join(EX, rel)
    if not os.path.exists(src):
        print("MISSING", rel); continue
    wd = tempfile.mkdtemp(prefix="sweep_")
    xml = open(src).read()
    # shrink: iteration counts to a handful, lattice as is (examples are small) but cap the big ones
    xml = re.sub(r'Iterations="(\d+)"', lambda m: 'Iterations="%d"' % max(1, min(int(m.group(1)), 6)), xml)
    open(os.path.join(wd, "case.xml"), "w").write(xml)
    for extra in glob.glob(os.path.join(os.path.dirname(src), "*")):
        if not extra.endswith(".xml") and os.path.isfile(extra) and os.path.getsize(extra) < 5e6:
            shutil.copy(extra, wd)
    try:
        r = subprocess.run([os.path.join(ROOT, "bin", "tclb"), "--cpu", "--precision", "double", model, "case.xml"], cwd=wd, capture_output=True, text=True, timeout=180)
        ok = r.returncode == 0
        msg = "" if ok else [l for l in (r.stdout + r.stderr).splitlines() if " E" in l[:8] or "FATAL" in l or "rror" in l][-3:]
    except subprocess.TimeoutExpired:
        ok, msg = False, "timeout (lattice too large for the CPU sweep)"
    print(("OK  " if ok else "FAIL"), model, rel, msg)
    bad += 0 if ok else 1
    shutil.rmtree(wd, ignore_errors=True)
print(bad, "failures")
