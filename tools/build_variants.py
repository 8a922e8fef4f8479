#!/usr/bin/env python
"""Build tuning variants of a model library under another name (diagnostics for gpurun experiments).

  python tools/build_variants.py d3q27_cumulant pw -DTCLB_GL_PER_WARP=1
builds tclb_b200/_lib/libtclb_d3q27_cumulant_pw_float.so, usable as `bench.py --model d3q27_cumulant_pw`.
"""
import sys

sys.path.insert(0, ".")
from tclb_b200 import build
from tclb_b200.models import build_model

model, tag, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
md = build_model(model)
print(build.build_custom(md, f"{model}_{tag}", "float", extra_flags=flags))
