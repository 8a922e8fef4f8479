#!/usr/bin/env python
"""Run a set of the reference's example cases (example/**.xml) for a few iterations on the CPU backend and report which
ones run: a smoke test of the XML front-end, geometry, handlers and units against real case files.  Edit `cases` to taste.
    python tools/sweep_examples.py
Known and expected failures: examples written for node types their model no longer declares (`Inlet`/`Outlet`/`XYslice1`
in flow/3d, `MRT` in electrokinetic/poison_boltzmann.xml), cases that need their MPMD partner program or files addressed
relative to the reference's root."""
import os, re, subprocess, sys, tempfile, shutil
ROOT="/root/repo"; EX="/root/reference/example"
BUILD = {   # library -> (model directory under models/, options)
 "ref_d2q9_pf_velocity": ("multiphase/d2q9_pf_velocity", None), "ref_d2q9q9_cm_cht": ("heat/d2q9q9_cm_cht", None),
 "ref_d3q27_viscoplastic": ("flow/d3q27_viscoplastic", None), "ref_sw": ("flow/sw", None),
 "ref_d2q9_reaction_diffusion_system_AllenCahn": ("reaction/d2q9_reaction_diffusion_system", {"AllenCahn": True}),
 "ref_d2q9_ShanChen": ("multiphase/d2q9_ShanChen", None),
}
cases = [
 ("ref_d2q9_pf_velocity", "multiphase/d2q9_pf_velocity/Droplet_Falling_2D.xml"), ("ref_d2q9_pf_velocity", "multiphase/d2q9_pf_velocity/RayleighTaylorInstability.xml"),
 ("ref_d2q9_pf_velocity", "multiphase/d2q9_pf_velocity/cm_dam_break.xml"), ("ref_d2q9_pf_velocity", "multiphase/d2q9_pf_velocity/mrt_dam_break.xml"),
 ("ref_d2q9q9_cm_cht", "heat/d2q9q9_Hot_Karman_Re100_Pr5.xml"), ("ref_d2q9q9_cm_cht", "heat/d2q9q9_lid_driven_cavity.xml"),
 ("ref_d3q27_viscoplastic", "nonnewtonian/viscoplastic_sphere_test.xml"), ("ref_d3q27_viscoplastic", "nonnewtonian/viscoplastic_tunnel_test.xml"),
 ("ref_sw", "shallowwater/" + (os.listdir(EX + "/shallowwater")[0] if os.path.isdir(EX + "/shallowwater") else "x.xml")),
 ("ref_d2q9_reaction_diffusion_system_AllenCahn", "reaction/d2q9_reaction_diffusion_system/AC_TRT_example.xml"),
 ("ref_d2q9_ShanChen", "multiphase/SchanChen/d2q9_sc.xml"),
]
sys.path.insert(0, ROOT)
from tclb_b200 import build as _build
import glob as _glob
for lib, (mdir, opts) in BUILD.items():
    found = _glob.glob("/root/reference/models/" + mdir) or _glob.glob("/root/reference/models/*/" + mdir.split("/")[-1]) or _glob.glob("/root/reference/models/*/*/" + mdir.split("/")[-1])
    try:
        out = _build.build_from_dir(found[0], precision="double", options=opts) if opts else _build.build_from_dir(found[0], precision="double")
        print("built", os.path.basename(str(out)))
    except Exception as e:
        print("BUILD FAILED", lib, str(e)[-300:])
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
