"""The reference's OWN generator code (src/conf.R and src/*.Rt, unmodified) executed by this repository's R-subset
interpreter + template engine - the same path that builds the reference arm of bench.py (baseline/build_reference.py).

What is pinned here:
  * the node-type packing and the table sizes conf.R derives equal the ones tclb_b200.gen.modeldef derives;
  * the 27 margin-block sizes of a checkpoint image (`FTabs::PreAlloc` of the expanded LatticeContainer.inc.cpp, i.e.
    the reference's MarginSizes algebra, src/conf.R:897-999) equal what csrc/core/pri.cpp writes, block by block, for
    2-D and 3-D models, stencil-1 and stencil-2 fields, 1 and 4 ranks - the `.pri` layout is checked against the
    reference's algebra, not only by round trip."""
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
REF = "/root/reference"

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not mounted")


def _blocks_from_reference(model, on=()):
    import build_reference as br
    t = br.expand_one(model, "LatticeContainer.inc.cpp.Rt", on)
    i = t.index("void FTabs::PreAlloc")
    body = t[i:t.index("\n}", i)]
    return {int(b[5:]): e for e, b in re.findall(r"size = \(size_t\) (.*?)\*sizeof\(storage_t\);\s*CudaPreAlloc\( \(void\*\*\)&(block\d+)", body)}


CASES = [("d2q9", (), "d2q9", (64, 40, 1)), ("d3q27_cumulant", ("AVG",), "d3q27_cumulant_AVG", (32, 24, 20)),
         ("d3q27_pf_velocity", (), "ref_d3q27_pf_velocity", (32, 16, 24)), ("d3q19_heat", (), "d3q19_heat", (32, 12, 16)),
         ("d2q9_poison_boltzmann", (), "d2q9_poison_boltzmann", (32, 48, 1))]


@pytest.mark.parametrize("ref_model,on,lib,shape", CASES)
@pytest.mark.parametrize("nranks", [1, 4])
def test_pri_block_sizes_equal_reference_margin_algebra(built, ref_model, on, lib, shape, nranks):
    from tclb_b200.runtime import Domain, Model
    if lib.startswith("ref_") and not os.path.exists(os.path.join(ROOT, "tclb_b200", "_lib", f"libtclb_{lib}_double.so")):
        pytest.skip(f"{lib} not built")
    expr = _blocks_from_reference(ref_model, on)
    m = Model(lib, "double")
    d = Domain(m, shape, "cpu", nranks=nranks)
    for r in range(nranks):
        _, _, _, nx, ny, nz = d.boxes[r]
        mine = d.pri_blocks(r)
        for b in range(1, 28):
            want = int(eval(expr[b], {"nx": nx, "ny": ny, "nz": nz})) if b in expr else 0
            assert mine[b - 1] == want, (ref_model, "rank", r, "block", b, mine[b - 1], expr.get(b))
    d.close()


def test_conf_r_tables_equal_modeldef():
    import build_reference as br
    from tclb_b200.models import build_model
    t = br.expand_one("d3q27_cumulant", "Consts.h.Rt")
    defs = dict(re.findall(r"#define\s+(\w+)\s+(0x[0-9a-fA-F]+|\d+)\s", t))
    m = build_model("d3q27_cumulant")
    m.finalize()
    for n in m.nodetypes:
        key = ("ZONE_" if n.group == "SETTINGZONE" else "NODE_") + n.name
        assert int(defs[key], 0) == n.value, (key, defs[key], n.value)
    for g in m.nodetypegroups:
        assert int(defs["NODE_" + g.name], 0) == g.mask, g.name
    assert int(defs["ZONE_SHIFT"]) == m.zone_shift and int(defs["ZONE_MAX"]) == m.zone_max
    assert int(defs["FIELDS"]) == len(m.fields) and int(defs["GLOBALS"]) == len(m.globals) and int(defs["SUM_GLOBALS"]) == m.sum_globals
    assert int(defs["SETTINGS"]) == len(m.settings) and int(defs["ZONESETTINGS"]) == len(m.zonesettings)
    for i, g in enumerate(m.globals):
        assert int(defs["GLOBALS_" + g.name]) == i
    for i, s_ in enumerate(m.settings):
        assert int(defs["SETTINGS_" + s_.name]) == i
    assert int(defs["IN_OBJ_OFFSET"]) == m.in_obj_offset
