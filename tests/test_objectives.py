"""Model objectives = expressions of the Globals with generated value + gradient functions (reference AddObjective,
src/conf.R:349-372, OF_<name> in src/Lists.cpp.Rt:17-34, <Objective> src/Handlers/acObjective.cpp)."""
import os

import numpy as np

from tclb_b200 import build
from tclb_b200.runtime import Model


def test_reference_plate_model_declares_efficiency_objectives(built, reference_dir):
    build.build_from_dir(os.path.join(reference_dir, "models/moving/d2q9_plate"), precision="double")
    m = Model("ref_d2q9_plate", "double")
    assert "EfficiencyX" in m.objectives and "EfficiencyY" in m.objectives and set(m.globals) <= set(m.objectives)
    vals = {g: 0.0 for g in m.globals}
    vals.update(ForceX=3.0, ForceY=-1.5, Power=0.25)
    v, grad = m.objective("EfficiencyX", vals)              # ForceX / Power
    assert abs(v - 12.0) < 1e-14
    want = np.zeros(len(m.globals))
    want[m.globals.index("ForceX")] = 1 / 0.25
    want[m.globals.index("Power")] = -3.0 / 0.25 ** 2
    assert np.allclose(grad, want, rtol=1e-14, atol=0)
    v, grad = m.objective("Power", vals)                    # every Global is an objective of its own
    assert v == 0.25 and grad[m.globals.index("Power")] == 1.0 and np.count_nonzero(grad) == 1


def test_objective_element_linearises_at_the_current_globals(built, tclb_bin, tmp_path):
    """<Objective FluxRatio="2"/> after a run: the InObj weights are 2 * d(OutletFlux/InletFlux)/d(Global) at the Globals of the
    last iteration and the Objective global is 2 * OutletFlux / InletFlux"""
    import csv
    import subprocess
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from golden_karman import KARMAN_XML
    from tclb_b200.gen.poly import Poly
    from tclb_b200.models import build_model
    md = build_model("d2q9")
    md.AddObjective("FluxRatio", Poly.sym("OutletFlux") * Poly.sym("InletFlux") ** -1)
    build.build_custom(md, "d2q9_fluxratio", "double")
    xml = KARMAN_XML.replace('<Solve Iterations="1000"', '<Solve Iterations="60"').replace("</CLBConfig>", '<Objective FluxRatio="2"/>\n<Log Iterations="1"/>\n<Solve Iterations="1"/>\n</CLBConfig>')
    xml = xml.replace('<Log Iterations="100"/>', '<Log Iterations="60"/>')
    (tmp_path / "k.xml").write_text(xml)
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "d2q9_fluxratio", "k.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    rows = list(csv.reader(open(tmp_path / "output" / "k_Log_P00_00000000.csv")))
    hdr = [h.strip() for h in rows[0]]
    at60 = dict(zip(hdr, [v.strip() for v in rows[1]]))
    assert int(float(at60["Iteration"])) == 60
    o, i_ = float(at60["OutletFlux"]), float(at60["InletFlux"])
    import re
    w = {k: float(v) for k, v in re.findall(r"Setting (\w+InObj) in zone\s+\(-1\) to (\S+)", r.stdout)}
    assert abs(w["OutletFluxInObj"] - 2 / i_) <= 1e-12 * abs(2 / i_), (w, 2 / i_)
    assert abs(w["InletFluxInObj"] + 2 * o / i_ ** 2) <= 1e-12 * abs(2 * o / i_ ** 2)
    assert w["PressureLossInObj"] == 0.0
