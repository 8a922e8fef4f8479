"""Physics smoke tests of the shipped models on the CPU oracle (analytic solutions / invariants)."""
import numpy as np
import pytest


def _model(name, precision="double"):
    from tclb_b200 import build
    from tclb_b200.runtime import Model
    build.build_core()
    build.build_model_lib(name, precision)
    return Model(name, precision)


def _poiseuille(model_name, collision, shape, visc_name, force_name, iters, tol):
    from tclb_b200.runtime import Domain
    m = _model(model_name)
    nx, ny, nz = shape
    d = Domain(m, shape, "cpu", nranks=2)
    fl = np.full((nz, ny, nx), m.flag(collision), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall")
    fl[:, -1, :] = m.flag("Wall")
    d.set_flags(fl)
    d.set_setting(visc_name, 0.1)
    d.set_setting(force_name, 1e-5)
    d.init()
    d.iterate(iters)
    u = d.get_quantity("U")[nz // 2, :, nx // 3, 0]
    H = ny - 2
    y = np.arange(ny) - 0.5
    ana = 1e-5 / (2 * 0.1) * y * (H - y)
    err = np.abs(u[1:-1] - ana[1:-1]).max() / ana.max()
    assert err < tol, err


def test_d2q9_poiseuille():
    _poiseuille("d2q9", "MRT", (32, 34, 1), "Viscosity", "GravitationX", 6000, 6e-3)


def test_d2q9_srt_poiseuille():
    _poiseuille("d2q9_SRT", "BGK", (32, 34, 1), "nu", "GravitationX", 6000, 6e-3)


def test_d3q19_poiseuille():
    _poiseuille("d3q19", "MRT", (16, 18, 8), "nu", "ForceX", 3000, 8e-3)


def test_d3q27_poiseuille():
    _poiseuille("d3q27", "MRT", (16, 18, 8), "nu", "ForceX", 3000, 8e-3)


def test_d3q27_cumulant_poiseuille():
    _poiseuille("d3q27_cumulant", "MRT", (16, 18, 8), "nu", "ForceX", 3000, 8e-3)


def test_diffusion2d_steady_profile():
    from tclb_b200.runtime import Domain
    m = _model("diffusion2D")
    d = Domain(m, (16, 33, 1), "cpu", nranks=2)
    fl = np.zeros((1, 33, 16), dtype=np.uint32)
    fl[0, 0, :] = m.flag("Dirichlet") | (1 << m.zone_shift)
    fl[0, -1, :] = m.flag("Dirichlet") | (2 << m.zone_shift)
    d.set_flags(fl)
    d.set_setting("diff_coeff", 0.45)
    d.set_setting("Value", 0.0, zone=0)
    d.set_setting("Value", 1.0, zone=1)
    d.set_setting("Value", 3.0, zone=2)
    d.init()
    d.iterate(8000)
    phi = d.get_quantity("Phi")[0, :, 4]
    ana = 1.0 + 2.0 * np.arange(33) / 32.0
    assert np.abs(phi - ana).max() < 1e-6


def test_wave2d_and_advection_build_and_run():
    from tclb_b200.runtime import Domain
    for name in ("wave2D", "advection_diffusion2D", "advection_diffusion2D_fields"):
        m = _model(name)
        d = Domain(m, (32, 32, 1), "cpu")
        d.set_flags(np.zeros((1, 32, 32), dtype=np.uint32))
        for k, v in dict(Speed=0.3, Viscosity=0.01, diff_coeff=0.3, ux=0.1, uy=0.0, Value=1.0).items():
            if k in m.settings or k in m.zonesettings:
                d.set_setting(k, v)
        d.init()
        d.iterate(20)
        q = list(m.quantities)[0]
        assert np.isfinite(d.get_quantity(q)).all()


def test_d3q19_heat_heater_raises_temperature():
    from tclb_b200.runtime import Domain
    m = _model("d3q19_heat")
    shape = (16, 18, 8)
    d = Domain(m, shape, "cpu", nranks=2)
    fl = np.full((8, 18, 16), m.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall")
    fl[:, -1, :] = m.flag("Wall")
    fl[:, 1, :] |= m.flag("Heater") | (1 << m.zone_shift)
    d.set_flags(fl)
    d.set_setting("nu", 0.1)
    d.set_setting("FluidAlpha", 0.1)
    d.set_setting("Temperature", 1.0, zone=0)
    d.set_setting("Temperature", 2.0, zone=1)
    d.init()
    d.iterate(1500)
    T = d.get_quantity("T")[3, :, 5]
    assert abs(T[1] - 2.0) < 0.01 and T[1] > T[5] > T[12] >= 0.999
    assert abs(d.get_quantity("Rho").sum() - 16 * 18 * 8) < 1e-6


def test_poison_boltzmann_multistage_action():
    from tclb_b200.runtime import Domain
    m = _model("d2q9_poison_boltzmann")
    assert m.actions["Iteration"] == [0, 1, 2]
    d = Domain(m, (32, 34, 1), "cpu", nranks=2)
    fl = np.full((1, 34, 32), m.flag("Collision"), dtype=np.uint32)
    fl[0, 0, :] = m.flag("Wall")
    fl[0, -1, :] = m.flag("Wall")
    d.set_flags(fl)
    for k, v in dict(tau_psi=1.0, n_inf=1e-4, z=1, el=1, kb=1, T=1, epsilon=1, dt=1, psi_bc=1.0, psi0=0.0).items():
        d.set_setting(k, v)
    d.init()
    d.iterate(3000)
    psi = d.get_quantity("Psi")[0, :, 5]
    assert psi[1] > psi[8] > psi[16] > 0.9          # screened potential decays away from the charged walls
    assert abs(psi[1] - psi[-2]) < 1e-9              # symmetric
