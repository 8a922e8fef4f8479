#!/usr/bin/env python
"""nvcc-build (sm_100a, double) every reference model directory that lowers: the complement of tools/compile_reference_models.py,
which only checks the host syntax.  python tools/build_all_reference_models.py [--keep]  (about 20 minutes on 6 cores).
Libraries this run created are deleted again unless --keep is given: the whole set is about 130 MB, and the tree that travels
to the GPU box has a size limit."""
import sys, os, concurrent.futures as cf, traceback
sys.path.insert(0, '/root/repo')
from tclb_b200 import build
OPTS = {"d2q9_scmp": {"Kupershtokh": True}, "d2q9_reaction_diffusion_system": {"AllenCahn": True}, "d3q27_PSM": {"NEBB": True}}
dirs = []
for d, _, files in sorted(os.walk("/root/reference/models")):
    if "conf.mk" in files: dirs.append(d)
skip = {"d2q9_diff","d2q9_heat_adj","d2q9_kuper_adj","d2q9_lbmpy","d2q9_lee","d2q9_part","d2q9_pf_pressureEvolution","d2q9_pp_LBL","d2q9_pp_MCMP","d2q9_scmp","d3q19_adj","d3q19_heat_adj","d3q19_heat_adj_prop","d3q19_kuper","d3q27_csf"}
KEEP = "--keep" in sys.argv
before = set(p.name for p in build.LIBDIR.glob("libtclb_ref_*"))
def one(d):
    name = os.path.basename(d)
    if name in skip: return name, "skipped (does not lower)"
    try:
        o = OPTS.get(name)
        p = build.build_from_dir(d, precision="double", options=o) if o else build.build_from_dir(d, precision="double")
        if not KEEP and p.name not in before:
            for q in (p, p.with_suffix(p.suffix + ".stamp")):
                if q.exists(): q.unlink()
        return name, "ok"
    except Exception as e:
        msg = str(e)
        k = msg.find("error")
        return name, "FAILED " + msg[max(0,k-120):k+200].replace("\n"," ")
with cf.ThreadPoolExecutor(6) as ex:
    res = list(ex.map(one, dirs))
bad = [r for r in res if r[1].startswith("FAILED")]
for n, r in res:
    if r != "ok": print(n, r[:330])
print(len([r for r in res if r[1]=="ok"]), "built with nvcc;", len(bad), "failed")
