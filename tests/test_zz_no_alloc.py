"""The allocation-free scan entry point (tclb_scan_array_static), called through the C ABI.

Collected last on purpose (file name): it has to give the right answer both on a fresh process and after every
other GPU test has used the device and the library.
"""
import ctypes as C
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ELEMS = 67108864


def _lib():
    from tclb_b200.runtime import core
    lib = core()
    lib.tclb_scan_array.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
    lib.tclb_scan_array.restype = C.c_int
    lib.tclb_scan_last_error.restype = C.c_char_p
    lib.tclb_scan_array_static.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
    lib.tclb_scan_array_static.restype = C.c_int
    return lib


def _both(lib, torch, t):
    """(existing entry, new entry) results for device tensor t, each as 3 python floats."""
    ref = torch.full((3,), -7.0, dtype=torch.float64, device=t.device)
    new = torch.full((3,), -9.0, dtype=torch.float64, device=t.device)
    err = C.create_string_buffer(256)
    assert lib.tclb_scan_array(t.data_ptr(), t.numel(), t.element_size(), None, ref.data_ptr()) == 0, lib.tclb_scan_last_error()
    rc = lib.tclb_scan_array_static(t.data_ptr(), t.numel(), t.element_size(), None, new.data_ptr(), err, len(err))
    assert rc == 0, err.value
    torch.cuda.synchronize(t.device)
    return ref.cpu().tolist(), new.cpu().tolist()


def _check(lib, torch, t):
    ref, new = _both(lib, torch, t)
    h = t.detach().cpu().numpy().astype(np.float64)
    want = [float(np.isnan(h).sum()), float(np.nanmin(h)), float(np.nanmax(h))]
    assert new == ref            # same kernel, same reduction tree: the two entry points must agree exactly
    assert new[0] == want[0]
    assert math.isclose(new[1], want[1], rel_tol=0, abs_tol=0) and math.isclose(new[2], want[2], rel_tol=0, abs_tol=0)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_static_scan_matches_allocating_scan_two_calls_in_a_row(dtype):
    import torch
    lib = _lib()
    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda:0").manual_seed(1234)
    dt = getattr(torch, dtype)
    # first call: many blocks, some NaNs
    a = torch.randn(5_000_017, generator=g, device="cuda:0", dtype=dt)
    a[::100_003] = float("nan")
    a[777] = -40.5
    a[4_999_999] = 91.25
    _check(lib, torch, a)
    # second call straight after, much shorter: the array still holds the first call's partials beyond block 2
    b = torch.randn(9_001, generator=g, device="cuda:0", dtype=dt) * 3 + 1
    b[17] = float("nan")
    _check(lib, torch, b)
    # and a third, one element, on a non-default stream
    s = torch.cuda.Stream(device=0)
    one = torch.tensor([2.5], device="cuda:0", dtype=dt)
    out = torch.zeros(3, dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    assert lib.tclb_scan_array_static(one.data_ptr(), 1, one.element_size(), C.c_void_p(s.cuda_stream), out.data_ptr(), None, 0) == 0
    s.synchronize()
    assert out.cpu().tolist() == [0.0, 2.5, 2.5]


def test_static_scan_largest_size_and_rejections():
    import torch
    lib = _lib()
    torch.cuda.set_device(0)
    big = torch.ones(MAX_ELEMS, device="cuda:0", dtype=torch.float32)
    big[MAX_ELEMS - 1] = -3.0
    big[12345] = float("nan")
    _check(lib, torch, big)
    out = torch.zeros(3, dtype=torch.float64, device="cuda:0")
    err = C.create_string_buffer(256)
    # one element too many: refused before anything is launched (the pointer is never touched)
    rc = lib.tclb_scan_array_static(big.data_ptr(), MAX_ELEMS + 1, 4, None, out.data_ptr(), err, len(err))
    assert rc == -3 and str(MAX_ELEMS).encode() in err.value and b"exceeds" in err.value
    assert lib.tclb_scan_array_static(big.data_ptr(), 10, 2, None, out.data_ptr(), err, len(err)) == -2 and err.value
    assert lib.tclb_scan_array_static(None, 10, 4, None, out.data_ptr(), err, len(err)) == -1
    assert lib.tclb_scan_array_static(big.data_ptr(), 0, 4, None, out.data_ptr(), None, 0) == -1
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0.0, 0.0, 0.0]
    # still healthy after the refusals
    _check(lib, torch, big[:100_000])


def test_static_scan_second_device():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    lib = _lib()
    try:
        torch.cuda.set_device(1)
        t = torch.randn(1_000_003, device="cuda:1", dtype=torch.float32)
        t[5] = float("nan")
        _check(lib, torch, t)
    finally:
        torch.cuda.set_device(0)
