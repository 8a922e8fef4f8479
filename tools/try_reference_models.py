#!/usr/bin/env python
"""Try to load every model of a TCLB source tree through the R front-end + template engine; print a status table."""
import os
import sys
import traceback

sys.path.insert(0, ".")
from tclb_b200.gen import rfrontend, rtemplate  # noqa: E402

root = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
only = sys.argv[2:] 
ok = bad = 0
for d, _, files in sorted(os.walk(os.path.join(root, "models"))):
    if "conf.mk" not in files:
        continue
    name = os.path.basename(d)
    if only and name not in only:
        continue
    try:
        m, I = rfrontend.load_dynamics_r(d, reference_src=os.path.join(root, "src"), finalize=False)
    except Exception as e:
        print(f"{name:40s} Dynamics.R FAILED: {type(e).__name__}: {str(e)[:150]}")
        bad += 1
        continue
    try:
        if "Dynamics.c.Rt" in files:
            code = rtemplate.render(I, I.globals, os.path.join(d, "Dynamics.c.Rt"), [os.path.join(root, "src"), root])
        else:
            code = rtemplate._read(os.path.join(d, "Dynamics.c"))
        print(f"{name:40s} ok ({len(code.splitlines())} lines)")
        ok += 1
    except Exception as e:
        msg = str(e).replace("\n", " ")[:170]
        print(f"{name:40s} template FAILED: {type(e).__name__}: {msg}")
        if only:
            traceback.print_exc()
        bad += 1
print(ok, "ok,", bad, "failed")
