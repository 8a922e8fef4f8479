"""GPU tests of the persistent stage kernels (`k_stage_persist`): persistent blocks with plain loads and with the
shared-memory ring fed by bulk asynchronous copies must reproduce the one-block-per-tile kernel bit for bit
(same node code, same operands), Globals to rounding (different summation order)."""
import os

import numpy as np
import pytest

from helpers import run_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def libs(built):
    from tclb_b200 import build
    for m in ("d2q9", "d3q19", "d3q27_cumulant"):
        for p in ("float", "double"):
            build.build_model_lib(m, p)


def _with_kernel(mode, fn):
    old = os.environ.get("TCLB_KERNEL")
    os.environ["TCLB_KERNEL"] = mode
    try:
        return fn()
    finally:
        if old is None:
            del os.environ["TCLB_KERNEL"]
        else:
            os.environ["TCLB_KERNEL"] = old


@pytest.mark.parametrize("model,shape,precision", [("d2q9", (256, 48, 1), "float"), ("d3q19", (128, 24, 20), "float"),
                                                   ("d3q27_cumulant", (256, 20, 12), "float"), ("d3q27_cumulant", (128, 16, 12), "double")])
@pytest.mark.parametrize("nranks", [1, 4])
def test_persistent_kernels_equal_classic(model, shape, precision, nranks):
    ref = _with_kernel("classic", lambda: run_case(model, precision, "cuda", shape, 30, nranks=nranks))
    n = float(np.prod(shape))
    for mode in ("persist", "ring"):
        got = _with_kernel(mode, lambda: run_case(model, precision, "cuda", shape, 30, nranks=nranks))
        for q in ref:
            if q.startswith("_"):
                continue
            assert np.array_equal(ref[q], got[q]), (mode, model, q, np.abs(ref[q] - got[q]).max())
        for f, v in ref["_fields"].items():
            assert np.array_equal(v, got["_fields"][f]), (mode, model, f)
        eps = 1e-12 if precision == "double" else 2e-7
        for k, v in ref["_globals"].items():
            assert abs(v - got["_globals"][k]) <= eps * (abs(v) + n), (mode, k, v, got["_globals"][k])


def test_persistent_globals_every_iteration_reproducible():
    """Globals accumulated over all tiles of a persistent block and folded once: bitwise reproducible run to run"""
    from helpers import setup_case
    from tclb_b200.runtime import Domain, Model

    def run():
        m = Model("d3q27_cumulant", "float")
        d = Domain(m, (256, 24, 16), "cuda", nranks=1)
        setup_case(d, m, (256, 24, 16))
        d.iterate(15, globals_every=True)
        g = d.globals()
        d.close()
        return g
    a = _with_kernel("ring", run)
    b = _with_kernel("ring", run)
    c = _with_kernel("classic", run)
    assert a == b
    for k in a:
        assert abs(a[k] - c[k]) <= 2e-6 * (abs(c[k]) + 256 * 24 * 16), (k, a[k], c[k])
