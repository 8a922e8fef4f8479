"""Unit system binding (native implementation in csrc/host/units.cpp)."""
from __future__ import annotations

import ctypes as C

from ..runtime import NativeError, core


class UnitEnv:
    def __init__(self):
        lib = core()
        lib.tclb_units_create.restype = C.c_void_p
        lib.tclb_units_destroy.argtypes = [C.c_void_p]
        lib.tclb_units_set.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p]
        lib.tclb_units_make_gauge.argtypes = [C.c_void_p]
        lib.tclb_units_alt.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]
        lib.tclb_units_si.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.lib = lib
        self.h = lib.tclb_units_create()

    def __del__(self):
        try:
            self.lib.tclb_units_destroy(self.h)
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise NativeError(self.lib.tclb_last_error().decode())

    def set_unit(self, name: str, value: str, gauge: str = "1"):
        self._chk(self.lib.tclb_units_set(self.h, name.encode(), str(value).encode(), str(gauge).encode()))

    def make_gauge(self):
        self._chk(self.lib.tclb_units_make_gauge(self.h))

    def alt(self, expr) -> float:
        out = C.c_double()
        self._chk(self.lib.tclb_units_alt(self.h, str(expr).encode(), C.byref(out)))
        return out.value

    def si(self, expr):
        out = C.c_double()
        ex = (C.c_double * 9)()
        self._chk(self.lib.tclb_units_si(self.h, str(expr).encode(), C.byref(out), ex))
        return out.value, list(ex)
