"""Lattice storage types (reference ./configure --with-storage=half|half-shift|float-shift): the lattice may be kept
narrower than the node code computes in, optionally as deviation from a per-field constant (AddDensity(shift=))."""
import numpy as np
import pytest

from helpers import run_case

SHAPE = (64, 40, 1)


@pytest.fixture(scope="module")
def ref():
    from tclb_b200 import build
    for p in ("double", "float", "double_float", "double_floatshift", "float_half", "float_halfshift"):
        build.build_model_lib("d2q9", p)
    return run_case("d2q9", "double", "cpu", SHAPE, 200)


def test_shifted_float_storage_under_double_arithmetic(ref):
    plain = run_case("d2q9", "double_float", "cpu", SHAPE, 200)
    shifted = run_case("d2q9", "double_floatshift", "cpu", SHAPE, 200)
    e_plain = np.abs(plain["U"] - ref["U"]).max()
    e_shift = np.abs(shifted["U"] - ref["U"]).max()
    assert e_shift < 1e-6 and e_shift < 0.6 * e_plain    # deviations from w_i keep more mantissa bits in 32-bit storage


def test_half_storage_with_shift_is_usable_and_plain_half_is_not(ref):
    hs = run_case("d2q9", "float_halfshift", "cpu", SHAPE, 200)
    h = run_case("d2q9", "float_half", "cpu", SHAPE, 200)
    umax = np.abs(ref["U"]).max()
    e_hs = np.abs(hs["U"] - ref["U"]).max()
    e_h = np.abs(h["U"] - ref["U"]).max()
    assert np.isfinite(hs["U"]).all() and e_hs < 0.05 * umax, (e_hs, umax)
    assert e_h > 3 * e_hs                                 # 11 significant bits of f ~ 0.1-0.4 swamp u ~ 1e-3


def test_half_storage_halves_the_memory_and_roundtrips_fields():
    from tclb_b200.runtime import Domain, Model
    m32, m16 = Model("d2q9", "float"), Model("d2q9", "float_halfshift")
    assert m16.bytes_per_update == 9 * 2 * 2 + 2 and m32.bytes_per_update == 9 * 4 * 2 + 2
    d = Domain(m16, SHAPE, "cpu")
    vals = (4.0 / 9.0 + 1e-3 * np.random.default_rng(1).standard_normal((1, 40, 64))).astype(np.float32)
    d.set_field("f[0]", vals)
    back = d.get_field("f[0]")
    assert np.abs(back - vals).max() < 2e-6               # stored as half of (f - 4/9): ~1e-3 * 2^-11


def test_half_storage_checkpoint_image_is_half_sized(tmp_path):
    from tclb_b200.runtime import Domain, Model
    sizes = {}
    for p in ("float", "float_halfshift"):
        d = Domain(Model("d2q9", p), SHAPE, "cpu")
        d.init()
        d.iterate(3)
        f = tmp_path / f"{p}.pri"
        d.save_pri(0, str(f))
        sizes[p] = f.stat().st_size
        d2 = Domain(Model("d2q9", p), SHAPE, "cpu")
        d2.load_pri(0, str(f))
        assert np.array_equal(d2.get_field("f[3]"), d.get_field("f[3]"))
    assert sizes["float_halfshift"] * 2 == sizes["float"]
