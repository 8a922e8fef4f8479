"""Adjoint beyond one stage on one subdomain (reference: multi-stage Action_*_Adj of src/Lattice.cu.Rt:542-613, the
adjoint scatter of stencil reads src/LatticeAccess.inc.cpp.Rt:349-373, snapshot levels :46-61,843-890):
  * reads through stencil accessors are differentiated (d2q9_kuper lowered from the reference's files: phi(dx,dy));
  * actions of several stages are swept back stage by stage (d2q9_poison_boltzmann: 3 stages);
  * several subdomains give the gradient of one (ghost fold into the neighbour's buffer);
  * a bounded tape (checkpoints + recomputation) gives the gradient of the full tape.
Every gradient is checked against central differences of the recorded objective along a random direction."""
import os

import numpy as np
import pytest

from tclb_b200 import build
from tclb_b200.runtime import Adjoint, Domain, Model


def _objective_sum(d, n):
    tot = 0.0
    for _ in range(n):
        d.iterate(1, globals_every=True)
        tot += d.globals()["Objective"]
    return tot


def _directional_check(d, m, n, fields, tol, frames=0, seed=3):
    base = {f: d.get_field(f).astype(np.float64) for f in m.fields}
    adj = Adjoint(d)
    if frames:
        adj.set_tape_frames(frames)
    adj.record(n)
    J0 = adj.objective
    held = adj.tape_frames()
    adj.reverse()
    grad = {f: adj.get_field(f).astype(np.float64) for f in fields}
    adj.close()
    rng = np.random.default_rng(seed)
    direction = {f: rng.standard_normal(base[f].shape) for f in fields}
    predicted = sum(float((grad[f] * direction[f]).sum()) for f in fields)
    eps = 1e-6

    def J(sign):
        for f in m.fields:
            d.set_field(f, base[f] + (sign * eps * direction[f] if f in direction else 0.0))
        return _objective_sum(d, n)

    fd = (J(+1) - J(-1)) / (2 * eps)
    assert J0 != 0.0 and predicted != 0.0
    assert abs(fd - predicted) <= tol * max(abs(fd), abs(predicted)), (fd, predicted)
    return grad, held


def _terminal_check(d, m, n, tol, seed=5, fields=None):
    """cost J = <w, state after n iterations>: w is seeded into the adjoint buffers before the backward sweep; `fields`
    restricts the weights and the perturbed direction (fields that hold flags or integer offsets are not differentiable)"""
    rng = np.random.default_rng(seed)
    base = {f: d.get_field(f).astype(np.float64) for f in m.fields}
    sel = set(fields) if fields is not None else set(m.fields)
    w = {f: rng.standard_normal(base[f].shape) * (1.0 if f in sel else 0.0) for f in m.fields}
    direction = {f: rng.standard_normal(base[f].shape) * (1.0 if f in sel else 0.0) for f in m.fields}
    adj = Adjoint(d)
    adj.record(n)
    for f in m.fields:
        adj.set_field(f, w[f])
    adj.reverse()
    predicted = sum(float((adj.get_field(f) * direction[f]).sum()) for f in m.fields)
    adj.close()
    eps = 1e-6

    def J(sign):
        for f in m.fields:
            d.set_field(f, base[f] + sign * eps * direction[f])
        d.iterate(n)
        return sum(float((w[f] * d.get_field(f)).sum()) for f in m.fields)

    fd = (J(+1) - J(-1)) / (2 * eps)
    assert predicted != 0.0 and abs(fd - predicted) <= tol * max(abs(fd), abs(predicted)), (fd, predicted)


def _channel(m, shape, backend="cpu", nranks=1):
    nx, ny, nz = shape
    d = Domain(m, shape, backend, nranks=nranks)
    fl = np.full((nz, ny, nx), m.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall")
    fl[:, -1, :] = m.flag("Wall")
    fl[:, ny // 3:ny // 3 + 3, nx // 4:nx // 4 + 3] = m.flag("Wall")
    d.set_flags(fl)
    return d


@pytest.mark.parametrize("nranks", [1, 2, 4])
def test_adjoint_on_several_subdomains(built, nranks):
    build.build_model_lib("d2q9", "double")
    build.build_adjoint_lib("d2q9", "double")
    m = Model("d2q9", "double")
    d = _channel(m, (24, 16, 1), nranks=nranks)
    d.set_setting("Viscosity", 0.05); d.set_setting("GravitationX", 1e-4)
    d.set_setting("OutletFluxInObj", 1.0); d.set_setting("PressureLossInObj", 0.3)
    fl = d.get_flags()
    fl[:, 1:-1, 3] |= m.flag("Inlet"); fl[:, 1:-1, -4] |= m.flag("Outlet")
    d.set_flags(fl)
    d.init(); d.iterate(10)
    grad, _ = _directional_check(d, m, 6, list(m.fields), 1e-6)
    d.close()
    if nranks > 1:
        ref = test_adjoint_on_several_subdomains.ref
        for f in grad:
            assert np.abs(grad[f] - ref[f]).max() <= 1e-13 * max(1.0, np.abs(ref[f]).max()), f
    else:
        test_adjoint_on_several_subdomains.ref = grad


def test_bounded_tape_recomputes_to_the_same_gradient(built):
    build.build_model_lib("d2q9", "double")
    build.build_adjoint_lib("d2q9", "double")
    m = Model("d2q9", "double")
    grads = []
    for frames in (0, 6):
        d = _channel(m, (24, 16, 1))
        d.set_setting("Viscosity", 0.05); d.set_setting("GravitationX", 1e-4)
        d.set_setting("OutletFluxInObj", 1.0)
        fl = d.get_flags(); fl[:, 1:-1, -4] |= m.flag("Outlet"); d.set_flags(fl)
        d.init(); d.iterate(5)
        adj = Adjoint(d)
        if frames:
            adj.set_tape_frames(frames)
        adj.record(53)
        held = adj.tape_frames()
        state_after = {f: d.get_field(f) for f in m.fields}
        adj.reverse()
        grads.append({f: adj.get_field(f) for f in m.fields})
        assert all(np.array_equal(d.get_field(f), state_after[f]) for f in m.fields)      # the lattice is put back where it was
        assert held == 53 if not frames else held <= 4
        adj.close(); d.close()
    for f in grads[0]:
        assert np.array_equal(grads[0][f], grads[1][f]), f          # recomputation repeats the same arithmetic


def test_multi_stage_action_adjoint(built):
    """3-stage Iteration (BaseIteration, CalcPsi, CalcSubiter): the model has no Globals, so the cost is a terminal one,
    J = <w, state after n iterations>, seeded into the adjoint before the sweep"""
    build.build_model_lib("d2q9_poison_boltzmann", "double")
    build.build_adjoint_lib("d2q9_poison_boltzmann", "double")
    m = Model("d2q9_poison_boltzmann", "double")
    assert len(m.actions["Iteration"]) == 3
    nx, ny, n = 16, 12, 4
    d = Domain(m, (nx, ny, 1), "cpu", nranks=1)
    first = [k for k, v in m.nodetypes.items() if v[0] != 0][0]
    fl = np.full((1, ny, nx), m.flag("BGK") if "BGK" in m.nodetypes else m.flag(first), dtype=np.uint32)
    d.set_flags(fl)
    d.init()
    rng = np.random.default_rng(5)
    for f in m.fields:
        a = d.get_field(f)
        d.set_field(f, a + 1e-2 * rng.standard_normal(a.shape))
    d.iterate(3)
    _terminal_check(d, m, n, 1e-6)
    d.close()


def test_stencil_reads_are_differentiated(built, reference_dir):
    build.build_from_dir(os.path.join(reference_dir, "models/multiphase/d2q9_kuper"), precision="double")
    build.build_adjoint_lib("ref_d2q9_kuper", "double")
    m = Model("ref_d2q9_kuper", "double")
    nx, ny = 20, 14
    d = Domain(m, (nx, ny, 1), "cpu", nranks=1)
    fl = np.full((1, ny, nx), m.flag("MRT") if "MRT" in m.nodetypes else m.flag("BGK"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall"); fl[:, -1, :] = m.flag("Wall")
    d.set_flags(fl)
    for k, v in (("omega", 1.0), ("nu", 0.1666), ("Density", 1.0), ("Temperature", 0.56), ("FAcc", 1.0), ("Magic", 0.01), ("MagicA", -0.152),
                 ("MagicF", -0.6666666), ("GravitationX", 1e-5), ("Wetting", 1.0)):
        if k in m.settings or k in m.zonesettings:
            d.set_setting(k, v)
    d.init()
    rng = np.random.default_rng(9)
    for f in m.fields:
        a = d.get_field(f)
        d.set_field(f, a * (1.0 + 1e-3 * rng.standard_normal(a.shape)))
    d.iterate(2)
    _terminal_check(d, m, 4, 1e-5)
    d.close()


@pytest.mark.gpu
def test_ten_thousand_iterations_inside_a_fixed_tape_budget_gpu():
    """unsteady adjoint over 10 000 iterations of a 256 x 256 lattice with at most 64 stored states (device memory of the
    tape: 64 frames instead of 10 000); the sweep recomputes between kept frames and leaves the lattice where it was"""
    import time
    build.build_model_lib("d2q9", "double")
    build.build_adjoint_lib("d2q9", "double")
    m = Model("d2q9", "double")
    d = _channel(m, (256, 256, 1), backend="cuda")
    d.set_setting("Viscosity", 0.05); d.set_setting("GravitationX", 1e-5)
    d.set_setting("OutletFluxInObj", 1.0)
    fl = d.get_flags(); fl[:, 1:-1, -4] |= m.flag("Outlet"); d.set_flags(fl)
    d.init(); d.iterate(10)
    adj = Adjoint(d)
    adj.set_tape_frames(64)
    t0 = time.time()
    adj.record(10000)
    held = adj.tape_frames()
    after = d.get_field(m.fields[1])
    adj.reverse()
    dt = time.time() - t0
    g = adj.get_field(m.fields[1])
    print("10000 iterations 256^2, 64-frame tape: %d frames held, record + reverse %.1f s, objective %.6g" % (held, dt, adj.objective))
    assert held <= 64 and np.isfinite(g).all() and np.abs(g).max() > 0 and adj.objective != 0.0
    assert np.array_equal(d.get_field(m.fields[1]), after)
    adj.close(); d.close()


def _pf_velocity_domain(m, shape, nranks):
    nx, ny, nz = shape
    d = Domain(m, shape, "cpu", nranks=nranks)
    fl = np.full((nz, ny, nx), m.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall"); fl[:, -1, :] = m.flag("Wall")
    d.set_flags(fl)
    for k, v in (("Density_h", 1.0), ("Density_l", 0.1), ("PhaseField_init", 1.0), ("sigma", 1e-3), ("M", 0.05), ("IntWidth", 3.0),
                 ("Viscosity_l", 0.16), ("Viscosity_h", 0.16), ("Radius", 4.0), ("CenterX", nx / 2), ("CenterY", ny / 2), ("CenterZ", nz / 2),
                 ("BubbleType", -1.0), ("GravitationY", -1e-5)):
        if k in m.settings or k in m.zonesettings:
            d.set_setting(k, v)
    d.init()
    return d


@pytest.mark.parametrize("nranks", [1, 2])
def test_pf_velocity_adjoint_against_finite_differences(built, reference_dir, nranks):
    """the reference's d3q27_pf_velocity (4-stage iteration, two lattices, 27 stencil reads of the phase field, integer
    temporaries in the model code) from its own files: terminal-cost adjoint against central differences on 1 and 2 subdomains"""
    build.build_from_dir(os.path.join(reference_dir, "models/multiphase/d3q27_pf_velocity"), precision="double")
    if build.build_adjoint_lib("ref_d3q27_pf_velocity", "double", check_only=True) is None and not os.environ.get("TCLB_SLOW_TESTS"):
        pytest.skip("the adjoint library of d3q27_pf_velocity is missing or does not match what the generator emits now (12 minutes of nvcc): TCLB_SLOW_TESTS=1 builds it here")
    build.build_adjoint_lib("ref_d3q27_pf_velocity", "double")
    m = Model("ref_d3q27_pf_velocity", "double")
    assert len(m.actions["Iteration"]) == 4
    d = _pf_velocity_domain(m, (12, 10, 8), nranks)
    d.iterate(2)
    # the populations of both lattices, the phase field and the velocity; not the wall-normal offsets and boundary marks the
    # model also keeps in fields (`nw_x`, `IsBoundary`, ...: integers in disguise, compared with == in the model code)
    phys = [f for f in m.fields if f[0] in "gh" and f[1:].isdigit() or f in ("PhaseF", "U", "V", "W", "pnorm")]
    assert len(phys) >= 27 + 15 + 4
    _terminal_check(d, m, 3, 1e-7, fields=phys)
    d.close()


def test_tape_spills_to_disk_and_gives_the_same_gradient(built, tmp_path, monkeypatch):
    """TCLB_TAPE_HOST_BYTES bounds the host part of the tape; the frames beyond it are files under TCLB_TAPE_DIR (the
    reference's snapshot levels end on disk too, src/Lattice.cu.Rt:200-233): same gradient, files removed afterwards"""
    build.build_model_lib("d2q9", "double")
    build.build_adjoint_lib("d2q9", "double")
    m = Model("d2q9", "double")

    def gradient():
        d = _channel(m, (24, 16, 1))
        d.set_setting("Viscosity", 0.05); d.set_setting("GravitationX", 1e-4)
        d.set_setting("OutletFluxInObj", 1.0)
        fl = d.get_flags(); fl[:, 1:-1, -4] |= m.flag("Outlet"); d.set_flags(fl)
        d.init(); d.iterate(5)
        adj = Adjoint(d)
        adj.record(20)
        seen = sorted(os.listdir(tmp_path))
        adj.reverse()
        g = {f: adj.get_field(f) for f in m.fields}
        adj.close(); d.close()
        return g, seen

    monkeypatch.setenv("TCLB_TAPE_DIR", str(tmp_path))
    in_memory, seen = gradient()
    assert seen == []
    monkeypatch.setenv("TCLB_TAPE_HOST_BYTES", str(3 * 24 * 18 * 9 * 8))      # room for about three frames in memory
    spilled, seen = gradient()
    assert len(seen) >= 15 and all(n.startswith("tclb_tape_") for n in seen)
    assert sorted(os.listdir(tmp_path)) == []                                  # every file removed when its frame was released
    for f in in_memory:
        assert np.array_equal(in_memory[f], spilled[f]), f
