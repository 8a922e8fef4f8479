#!/usr/bin/env python
"""Build tuning variants of a model library under another name (diagnostics for gpurun experiments).

  python tools/build_variants.py d3q27_cumulant pw -DTCLB_GL_PER_WARP=1
      -> tclb_b200/_lib/libtclb_d3q27_cumulant_pw_float.so, usable as `bench.py --model d3q27_cumulant_pw`
  python tools/build_variants.py dir:/root/reference/models/multiphase/d3q27_pf_velocity mb5 -DTCLB_MIN_BLOCKS=5
      -> libtclb_ref_d3q27_pf_velocity_mb5_float.so from an unmodified TCLB model directory
"""
import os
import sys

sys.path.insert(0, ".")
from tclb_b200 import build
from tclb_b200.models import build_model

model, tag, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
if model.startswith("dir:"):
    from tclb_b200.gen import rfrontend
    md = rfrontend.load_model(model[4:])
    name = "ref_" + md.name
else:
    md = build_model(model)
    name = model
print(build.build_custom(md, f"{name}_{tag}", "float", extra_flags=flags))
