"""The hand-written native models against the SAME models lowered from the reference's unmodified model directories
(R front-end + template engine): fields after 50 iterations agree to rounding and the Globals agree, on a case with walls,
a body force, an obstacle, a Zou-He velocity inlet and a pressure outlet.  This pins the physics of the benchmarked
native d3q27_cumulant (central-moment arrangement) and d3q19 to the reference's own node code."""
import os

import numpy as np
import pytest

from tclb_b200 import build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(name, shape=(24, 16, 12), iters=50):
    from tclb_b200.runtime import Domain, Model
    m = Model(name, "double")
    d = Domain(m, shape, "cpu", nranks=1)
    nx, ny, nz = shape
    fl = np.full((nz, ny, nx), m.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall")
    fl[:, -1, :] = m.flag("Wall")
    fl[:, 1:-1, 0] = m.flag("WVelocity") | m.flag("MRT")
    fl[:, 1:-1, -1] = m.flag("EPressure") | m.flag("MRT")
    fl[4:7, 6:9, 8:11] = m.flag("Wall")
    d.set_flags(fl)
    for k, v in (("nu", 0.04), ("ForceX", 2e-5), ("ForceY", -1e-5), ("Velocity", 0.02), ("InletVelocity", 0.02)):
        if k in m.settings or k in m.zonesettings:
            d.set_setting(k, v)
    d.init()
    d.iterate(iters, globals_on_last=True)
    out = {f: d.get_field(f) for f in m.fields}, d.globals(), {q: d.get_quantity(q) for q in m.quantities}, list(m.fields)
    d.close()
    return out


@pytest.mark.parametrize("native,lowered,ref_dir", [("d3q27_cumulant", "ref_d3q27_cumulant", "flow/d3q27_cumulant"), ("d3q19", "ref_d3q19", "flow/d3q19")])
def test_native_equals_reference_lowering(built, reference_dir, native, lowered, ref_dir):
    build.build_model_lib(native, "double")
    build.build_from_dir(os.path.join(reference_dir, "models", ref_dir), precision="double")
    fa, ga, qa, na = _run(native)
    fb, gb, qb, nb = _run(lowered)
    assert na == nb and list(ga) == list(gb) and list(qa) == list(qb)
    assert float(np.abs(qa["U"]).max()) > 0.02                  # the flow is not trivial
    for f in fa:
        assert np.abs(fa[f] - fb[f]).max() <= 1e-12, (native, f, np.abs(fa[f] - fb[f]).max())
    for q in qa:
        assert np.abs(qa[q] - qb[q]).max() <= 1e-12, (native, q)
    for k in ga:
        assert abs(ga[k] - gb[k]) <= 1e-10 * max(1.0, abs(ga[k])), (native, k, ga[k], gb[k])
