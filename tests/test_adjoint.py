"""Discrete adjoint (dual-number build of the model library + reverse sweep over a tape) against finite differences."""
import numpy as np
import pytest

from tclb_b200 import build
from tclb_b200.runtime import Adjoint, Domain, Model


def _channel(m, shape, backend):
    nx, ny, nz = shape
    d = Domain(m, shape, backend, nranks=1)
    fl = np.zeros((nz, ny, nx), dtype=np.uint32)
    fl[:] = m.nodetypes["MRT"][0]
    fl[:, 0, :] = m.nodetypes["Wall"][0]
    fl[:, -1, :] = m.nodetypes["Wall"][0]
    fl[:, 1:-1, 2] |= m.nodetypes["Inlet"][0]
    fl[:, 1:-1, nx - 3] |= m.nodetypes["Outlet"][0]
    d.set_flags(fl)
    d.set_setting("Viscosity", 0.05)
    d.set_setting("GravitationX", 1e-4)
    d.set_setting("OutletFluxInObj", 1.0)
    d.set_setting("PressureLossInObj", 0.5)
    d.init()
    d.iterate(20)
    return d


def _objective_sum(d, n):
    tot = 0.0
    for _ in range(n):
        d.iterate(1, globals_every=True)
        tot += d.globals()["Objective"]
    return tot


def _check_against_fd(backend, precision, eps, tol):
    build.build_model_lib("d2q9", precision)
    build.build_adjoint_lib("d2q9", precision)
    m = Model("d2q9", precision)
    shape = (24, 12, 1)
    n = 6
    d = _channel(m, shape, backend)
    fields = list(m.fields)
    base = {f: d.get_field(f).astype(np.float64) for f in fields}
    adj = Adjoint(d)
    adj.record(n)
    J0 = adj.objective
    adj.reverse()
    grad = {f: adj.get_field(f).astype(np.float64) for f in fields}
    assert J0 != 0.0 and any(np.abs(g).max() > 0 for g in grad.values())
    rng = np.random.default_rng(3)
    direction = {f: rng.standard_normal(base[f].shape) for f in fields}
    predicted = sum(float((grad[f] * direction[f]).sum()) for f in fields)

    def J(sign):
        for f in fields:
            d.set_field(f, base[f] + sign * eps * direction[f])
        return _objective_sum(d, n)

    fd = (J(+1) - J(-1)) / (2 * eps)
    assert abs(fd - predicted) <= tol * max(abs(fd), abs(predicted)), (fd, predicted)
    # the same run without the tape gives the same objective
    for f in fields:
        d.set_field(f, base[f])
    assert abs(_objective_sum(d, n) - J0) <= 1e-6 * abs(J0) + (1e-12 if precision == "double" else 1e-4)
    adj.close()
    d.close()


def test_adjoint_gradient_matches_finite_differences_cpu():
    _check_against_fd("cpu", "double", 1e-6, 1e-6)


def test_adjoint_needs_its_library(tmp_path):
    build.build_model_lib("d2q9_SRT", "double")
    m = Model("d2q9_SRT", "double")
    d = Domain(m, (16, 8, 1), "cpu", nranks=1)
    d.init()
    import os
    lib = build.LIBDIR / "libtclb_d2q9_SRT_double_ad.so"
    if not lib.exists():
        with pytest.raises(Exception, match="build_adjoint_lib"):
            Adjoint(d)
    d.close()


@pytest.mark.gpu
def test_adjoint_gradient_matches_finite_differences_gpu():
    _check_against_fd("cuda", "double", 1e-6, 1e-6)
    _check_against_fd("cuda", "float", 2e-3, 5e-2)


def _run_adj_case(reference_dir, tclb_bin, tmp_path, backend_flag, precision):
    """The reference's own topology-optimisation model (models/optimization/d2q9_adj, lowered by the front-end, primal
    part + dual-number adjoint) through <Objective>, <FDTest>, <Adjoint>, <VTK what=w_adj>, <OptSolve>, <ThresholdNow>."""
    import os
    import re
    import shutil
    import subprocess
    if reference_dir is not None:
        build.build_from_dir(os.path.join(reference_dir, "models/optimization/d2q9_adj"), precision=precision)
        build.build_adjoint_lib("ref_d2q9_adj", precision)
    elif not (build.LIBDIR / f"libtclb_ref_d2q9_adj_{precision}_ad.so").exists():
        pytest.skip("ref_d2q9_adj libraries were not built (reference tree not mounted at build time)")
    here = os.path.dirname(os.path.abspath(__file__))
    shutil.copy(os.path.join(here, "cases", "adj_drag_small.xml"), tmp_path / "adj.xml")
    r = subprocess.run([tclb_bin, backend_flag, "--precision", precision, "ref_d2q9_adj", "adj.xml"], cwd=tmp_path,
                       capture_output=True, text=True, timeout=600)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    worst = float(re.search(r"FDTest: largest relative difference ([-+.eE0-9]+)", out).group(1))
    sweep = re.search(r"adjoint swept back over (\d+) iterations, Objective ([-+.eE0-9]+)", out)
    fd_obj = float(re.search(r"FDTest: objective ([-+.eE0-9]+)", out).group(1))
    assert int(sweep.group(1)) == 8
    assert abs(float(sweep.group(2)) - fd_obj) <= 1e-6 * abs(fd_obj)      # same 8 iterations from the same state
    ev = [float(v) for v in re.findall(r"Optimize: evaluation \d+ objective ([-+.eE0-9]+)", out)]
    best = float(re.search(r"Optimize: best objective ([-+.eE0-9]+)", out).group(1))
    assert len(ev) >= 2 and best >= ev[0] and best == max(ev)              # maximised, never worse than the start
    assert "Parameters: 192" in out                                       # the 16x12 design space was thresholded
    vti = [f for f in os.listdir(tmp_path / "output") if f.endswith(".vti")]
    assert vti and b"w_adj" in open(tmp_path / "output" / vti[0], "rb").read()
    return worst


def test_reference_adjoint_model_fdtest_and_optsolve_cpu(reference_dir, tclb_bin, tmp_path):
    assert _run_adj_case(reference_dir, tclb_bin, tmp_path, "--cpu", "double") < 1e-5


@pytest.mark.gpu
def test_reference_adjoint_model_fdtest_and_optsolve_gpu(tclb_bin, tmp_path):
    # the libraries travel with the tree; the reference checkout itself is not on the GPU box
    assert _run_adj_case(None, tclb_bin, tmp_path, "--gpu", "double") < 1e-5


def test_adjoint_setting_gradients_match_finite_differences_cpu():
    """dObjective/d(setting) from the settings pass of the adjoint, including the derived chain Viscosity -> omega"""
    build.build_model_lib("d2q9", "double")
    build.build_adjoint_lib("d2q9", "double")
    m = Model("d2q9", "double")
    shape, n = (24, 12, 1), 6
    d = _channel(m, shape, "cpu")
    fields = list(m.fields)
    base = {f: d.get_field(f) for f in fields}
    names = ["Viscosity", "GravitationX"]
    values = {"Viscosity": 0.05, "GravitationX": 1e-4}
    adj = Adjoint(d)
    adj.want_setting_gradients(names)
    adj.record(n)
    adj.reverse()
    for name in names:
        g = adj.setting_gradient(name)
        eps = 1e-6 * max(1.0, abs(values[name]) * 1e3)
        J = []
        for sign in (+1, -1):
            d.set_setting(name, values[name] + sign * eps)
            for f in fields:
                d.set_field(f, base[f])
            J.append(_objective_sum(d, n))
        d.set_setting(name, values[name])
        fd = (J[0] - J[1]) / (2 * eps)
        assert g != 0.0 and abs(fd - g) <= 1e-5 * max(abs(fd), abs(g)), (name, fd, g)
    adj.close()
    d.close()


def test_adjoint_3d_periodic_fold_matches_finite_differences_cpu():
    """d3q19 on a box that is periodic in y and z: the contributions scattered into the y/z ghost layers (edges and
    corners included) have to come back through the ghost fold"""
    build.build_model_lib("d3q19", "double")
    build.build_adjoint_lib("d3q19", "double")
    m = Model("d3q19", "double")
    shape, n = (10, 8, 6), 5
    nx, ny, nz = shape
    d = Domain(m, shape, "cpu", nranks=1)
    fl = np.full((nz, ny, nx), m.nodetypes["MRT"][0], dtype=np.uint32)
    fl[2, 3, 4] = m.nodetypes["Wall"][0]                       # one obstacle, no walls: fully periodic
    d.set_flags(fl)
    d.set_setting("nu", 0.05)
    d.set_setting("ForceX", 1e-4)
    d.set_setting("ForceY", 5e-5)
    d.set_setting("ForceZ", -3e-5)
    d.set_setting("FluxInObj", 1.0)
    d.set_setting("VOLvyInObj", 0.7)
    d.set_setting("VOLpzInObj", -0.4)
    d.init()
    d.iterate(10)
    fields = list(m.fields)
    base = {f: d.get_field(f).astype(np.float64) for f in fields}
    adj = Adjoint(d)
    adj.record(n)
    J0 = adj.objective
    adj.reverse()
    grad = {f: adj.get_field(f).astype(np.float64) for f in fields}
    rng = np.random.default_rng(11)
    direction = {f: rng.standard_normal(base[f].shape) for f in fields}
    predicted = sum(float((grad[f] * direction[f]).sum()) for f in fields)
    eps = 1e-6

    def J(sign):
        for f in fields:
            d.set_field(f, base[f] + sign * eps * direction[f])
        return _objective_sum(d, n)

    fd = (J(+1) - J(-1)) / (2 * eps)
    assert J0 != 0.0 and abs(fd - predicted) <= 1e-6 * max(abs(fd), abs(predicted)), (fd, predicted)
    adj.close()
    d.close()


def test_recorded_segment_reports_time_integrated_globals_cpu():
    """Globals read after a recorded segment are integrated over its iterations, like the reference's ITER_INTEG
    (src/Lattice.h.Rt:29; calcGlobals adds the device totals of every iteration, src/Lattice.cu.Rt:1313)."""
    build.build_model_lib("d2q9", "double")
    build.build_adjoint_lib("d2q9", "double")
    m = Model("d2q9", "double")
    d = _channel(m, (24, 12, 1), "cpu")
    base = {f: d.get_field(f).astype(np.float64) for f in m.fields}
    n = 5
    sums = {}
    for _ in range(n):
        d.iterate(1, globals_every=True)
        for k, v in d.globals().items():
            sums[k] = sums.get(k, 0.0) + v
    for f in m.fields:
        d.set_field(f, base[f])
    adj = Adjoint(d)
    adj.record(2)
    adj.record(n - 2)              # two calls inside one tape keep integrating
    got = d.globals()
    for k in ("Objective", "OutletFlux", "PressureLoss"):
        assert abs(got[k] - sums[k]) <= 1e-10 * max(1.0, abs(sums[k])), (k, got[k], sums[k])
    assert abs(adj.objective - sums["Objective"]) <= 1e-10 * max(1.0, abs(sums["Objective"]))
    adj.close()
    d.close()
