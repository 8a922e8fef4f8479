"""Option autosym (reference src/conf.R:373-457, SymmetryAccess in src/LatticeAccess.inc.cpp.Rt:376-475; variants
<model>_autosym / _autosym2, src/models.R:52-57): a half channel closed by a symmetry plane reproduces the lower half of the
full channel - plane through the top node row (autosym, Symmetry*) and plane on the cell edge above it (autosym2,
SymmetryEdge*), for the native d2q9 and for the library lowered from the reference's own d2q9 files."""
import os

import numpy as np
import pytest

from tclb_b200 import build


def _run(lib, ny, sym_flag, iters=300, nx=32):
    from tclb_b200.runtime import Domain, Model
    m = Model(lib, "double")
    d = Domain(m, (nx, ny, 1), "cpu", nranks=1)
    fl = np.full((1, ny, nx), m.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall")
    if sym_flag:
        fl[:, -1, :] |= m.flag(sym_flag)
    else:
        fl[:, -1, :] = m.flag("Wall")
    fl[0, 3:5, 10:13] = m.flag("Wall")              # an obstacle near the wall, mirrored in the full channel below
    if not sym_flag:
        fl[0, ny - 5:ny - 3, 10:13] = m.flag("Wall")
    d.set_flags(fl)
    d.set_setting("Viscosity", 0.05)
    d.set_setting("GravitationX", 2e-5)
    d.init()
    d.iterate(iters)
    u = d.get_quantity("U")
    rho = d.get_quantity("Rho")
    d.close()
    return u[0], rho[0]


@pytest.mark.parametrize("lib,from_reference", [("d2q9", False), ("ref_d2q9", True)])
@pytest.mark.parametrize("mode", [1, 2])
def test_half_channel_with_symmetry_plane_equals_full_channel(built, reference_dir, lib, from_reference, mode):
    H = 12
    if from_reference:
        build.build_from_dir(os.path.join(reference_dir, "models/flow/d2q9"), precision="double")
        build.build_from_dir(os.path.join(reference_dir, "models/flow/d2q9"), precision="double", options={"autosym": mode})
        sym_lib = "ref_d2q9_autosym" + ("2" if mode == 2 else "")
    else:
        build.build_model_lib("d2q9", "double")
        build.build_model_lib("d2q9_autosym" + ("2" if mode == 2 else ""), "double")
        sym_lib = "d2q9_autosym" + ("2" if mode == 2 else "")
    ny_full = 2 * H - 1 if mode == 1 else 2 * H
    name = ("Symmetry" if mode == 1 else "SymmetryEdge") + "Y_plus"
    u_full, rho_full = _run(lib, ny_full, None)
    u_half, rho_half = _run(sym_lib, H, name)
    assert float(np.abs(u_full[..., 0]).max()) > 1e-3
    # the full channel is symmetric itself (to rounding), and its lower half is what the half channel computes
    # (row 0 is the wall row: what a wall node pulls from beyond the periodic wrap has no meaning)
    assert np.abs(u_full[1:H] - u_half[1:]).max() < 1e-12, np.abs(u_full[1:H] - u_half[1:]).max()
    assert np.abs(rho_full[1:H] - rho_half[1:]).max() < 1e-12
    assert np.abs(u_full[::-1][1:H, :, 0] - u_full[1:H, :, 0]).max() < 1e-12
