"""ctypes binding of the native runtime (``tclb_b200/_lib/libtclb_core.so``).

The Python layer never re-implements the engine: every call below lands in the C ABI declared in
``include/tclb_b200.h``.  On a GPU box a missing native library is an error, never a fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Dict, List, Optional, Sequence

import numpy as np

ROOT = Path(__file__).resolve().parent
LIBDIR = ROOT / "_lib"

BACKEND_CPU, BACKEND_CUDA = 0, 1
GLOBALS_NONE, GLOBALS_INTEGRATE, GLOBALS_OBJECTIVE = 0, 1, 2


class _Density(C.Structure):
    _fields_ = [("name", C.c_char_p), ("field", C.c_char_p), ("group", C.c_char_p), ("comment", C.c_char_p),
                ("dx", C.c_int), ("dy", C.c_int), ("dz", C.c_int), ("field_index", C.c_int),
                ("parameter", C.c_int), ("average", C.c_int), ("adjoint", C.c_int), ("has_default", C.c_int),
                ("default_value", C.c_double)]


class _Field(C.Structure):
    _fields_ = [("name", C.c_char_p), ("group", C.c_char_p), ("comment", C.c_char_p),
                ("minx", C.c_int), ("maxx", C.c_int), ("miny", C.c_int), ("maxy", C.c_int), ("minz", C.c_int), ("maxz", C.c_int),
                ("parameter", C.c_int), ("average", C.c_int), ("adjoint", C.c_int)]


class _Setting(C.Structure):
    _fields_ = [("name", C.c_char_p), ("unit", C.c_char_p), ("default_value", C.c_char_p), ("comment", C.c_char_p),
                ("zonal", C.c_int), ("derived_target", C.c_int), ("derive", C.c_void_p), ("adjoint", C.c_int)]


class _Global(C.Structure):
    _fields_ = [("name", C.c_char_p), ("unit", C.c_char_p), ("comment", C.c_char_p), ("op", C.c_int), ("adjoint", C.c_int),
                ("base", C.c_double)]


class _Quantity(C.Structure):
    _fields_ = [("name", C.c_char_p), ("unit", C.c_char_p), ("comment", C.c_char_p), ("vector", C.c_int), ("adjoint", C.c_int)]


class _NodeType(C.Structure):
    _fields_ = [("name", C.c_char_p), ("group", C.c_char_p), ("value", C.c_uint32), ("mask", C.c_uint32),
                ("shift", C.c_int), ("bits", C.c_int), ("save", C.c_int)]


class _NodeGroup(C.Structure):
    _fields_ = [("name", C.c_char_p), ("mask", C.c_uint32), ("shift", C.c_int), ("bits", C.c_int), ("capacity", C.c_int),
                ("save", C.c_int)]


class _Stage(C.Structure):
    _fields_ = [("name", C.c_char_p), ("main", C.c_char_p), ("particle", C.c_int),
                ("load_density", C.POINTER(C.c_ubyte)), ("save_field", C.POINTER(C.c_ubyte)), ("read_field", C.POINTER(C.c_ubyte))]


class _Action(C.Structure):
    _fields_ = [("name", C.c_char_p), ("n_stages", C.c_int), ("stages", C.POINTER(C.c_int))]


class _Option(C.Structure):
    _fields_ = [("name", C.c_char_p), ("value", C.c_int)]


class _Objective(C.Structure):
    _fields_ = [("name", C.c_char_p), ("fun", C.CFUNCTYPE(None, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)))]


class _ModelDesc(C.Structure):
    _fields_ = [("abi_version", C.c_int), ("name", C.c_char_p), ("base_name", C.c_char_p),
                ("real_size", C.c_int), ("flag_size", C.c_int),
                ("zone_shift", C.c_int), ("zone_max", C.c_int), ("sum_globals", C.c_int), ("in_obj_offset", C.c_int),
                ("n_densities", C.c_int), ("densities", C.POINTER(_Density)),
                ("n_fields", C.c_int), ("fields", C.POINTER(_Field)),
                ("n_settings", C.c_int), ("settings", C.POINTER(_Setting)),
                ("n_zonesettings", C.c_int), ("zonesettings", C.POINTER(_Setting)),
                ("n_globals", C.c_int), ("globals", C.POINTER(_Global)),
                ("n_quantities", C.c_int), ("quantities", C.POINTER(_Quantity)),
                ("n_nodetypes", C.c_int), ("nodetypes", C.POINTER(_NodeType)),
                ("n_nodegroups", C.c_int), ("nodegroups", C.POINTER(_NodeGroup)),
                ("n_stages", C.c_int), ("stages", C.POINTER(_Stage)),
                ("n_actions", C.c_int), ("actions", C.POINTER(_Action)),
                ("n_options", C.c_int), ("options", C.POINTER(_Option)),
                ("gy", C.c_int), ("gz", C.c_int),
                ("run_stage", C.c_void_p), ("get_quantity", C.c_void_p), ("color", C.c_void_p),
                ("kernel_info", C.c_void_p), ("last_error", C.c_void_p),
                ("storage_size", C.c_int), ("storage_shifted", C.c_int), ("storage_name", C.c_char_p),
                ("field_shift", C.POINTER(C.c_double)), ("part_margin", C.c_double),
                ("n_objectives", C.c_int), ("objectives", C.POINTER(_Objective))]


_core = None


class NativeError(RuntimeError):
    pass


def core_path() -> Path:
    return LIBDIR / "libtclb_core.so"


def core():
    """Load libtclb_core.so (building it first if it is missing and a compiler is available)."""
    global _core
    if _core is not None:
        return _core
    p = core_path()
    if not p.exists():
        from . import build
        build.build_core()
    lib = C.CDLL(str(p), mode=C.RTLD_GLOBAL)
    lib.tclb_last_error.restype = C.c_char_p
    lib.tclb_model_load.restype = C.c_void_p
    lib.tclb_model_load.argtypes = [C.c_char_p]
    lib.tclb_model_desc.restype = C.POINTER(_ModelDesc)
    lib.tclb_model_desc.argtypes = [C.c_void_p]
    lib.tclb_domain_create.restype = C.c_void_p
    lib.tclb_domain_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int), C.c_int]
    lib.tclb_domain_destroy.argtypes = [C.c_void_p]
    lib.tclb_domain_destroy.restype = None
    lib.tclb_adjoint_create.restype = C.c_void_p
    lib.tclb_adjoint_destroy.restype = None
    lib.tclb_adjoint_objective.restype = C.c_double
    lib.tclb_domain_get_setting.restype = C.c_double
    lib.tclb_domain_snapshot_ptr.restype = C.c_void_p
    lib.tclb_domain_flags_ptr.restype = C.c_void_p
    lib.tclb_domain_stream.restype = C.c_void_p
    for fn, args in {
        "tclb_domain_decomposition": [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)],
        "tclb_domain_set_halo_mode": [C.c_void_p, C.c_int],
        "tclb_domain_set_setting": [C.c_void_p, C.c_char_p, C.c_double],
        "tclb_domain_get_setting": [C.c_void_p, C.c_char_p],
        "tclb_domain_set_zone_setting": [C.c_void_p, C.c_char_p, C.c_int, C.c_double],
        "tclb_domain_set_zone_series": [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_double), C.c_int],
        "tclb_domain_set_flags": [C.c_void_p, C.c_int, C.c_void_p],
        "tclb_domain_get_flags": [C.c_void_p, C.c_int, C.c_void_p],
        "tclb_domain_set_cuts": [C.c_void_p, C.c_int, C.c_void_p],
        "tclb_domain_init": [C.c_void_p],
        "tclb_domain_iterate": [C.c_void_p, C.c_int, C.c_int, C.c_int],
        "tclb_domain_run_action": [C.c_void_p, C.c_char_p, C.c_int],
        "tclb_domain_sync": [C.c_void_p],
        "tclb_domain_iter": [C.c_void_p],
        "tclb_domain_get_globals": [C.c_void_p, C.POINTER(C.c_double)],
        "tclb_domain_get_field": [C.c_void_p, C.c_int, C.c_char_p, C.c_void_p],
        "tclb_domain_set_field": [C.c_void_p, C.c_int, C.c_char_p, C.c_void_p],
        "tclb_domain_get_quantity": [C.c_void_p, C.c_int, C.c_char_p, C.POINTER(C.c_int), C.c_void_p, C.c_double],
        "tclb_domain_color": [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p],
        "tclb_domain_reset_average": [C.c_void_p],
        "tclb_domain_set_block": [C.c_void_p, C.c_int, C.c_int],
        "tclb_domain_use_graphs": [C.c_void_p, C.c_int],
        "tclb_domain_graph_iterations": [C.c_void_p],
        "tclb_domain_save_pri": [C.c_void_p, C.c_int, C.c_char_p],
        "tclb_domain_pri_blocks": [C.c_void_p, C.c_int, C.POINTER(C.c_longlong)],
        "tclb_domain_load_pri": [C.c_void_p, C.c_int, C.c_char_p],
        "tclb_domain_snapshot_ptr": [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)],
        "tclb_domain_flags_ptr": [C.c_void_p, C.c_int],
        "tclb_domain_stream": [C.c_void_p, C.c_int],
        "tclb_domain_export_ipc": [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int64)],
        "tclb_domain_import_peer": [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int64), C.c_int, C.c_int],
        "tclb_domain_finalize_wiring": [C.c_void_p],
        "tclb_domain_set_hooks": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
        "tclb_adjoint_create": [C.c_void_p],
        "tclb_adjoint_destroy": [C.c_void_p],
        "tclb_adjoint_record": [C.c_void_p, C.c_int],
        "tclb_adjoint_reverse": [C.c_void_p],
        "tclb_adjoint_steady": [C.c_void_p, C.c_int],
        "tclb_adjoint_zero": [C.c_void_p],
        "tclb_adjoint_set_tape_frames": [C.c_void_p, C.c_int],
        "tclb_adjoint_tape_frames": [C.c_void_p],
        "tclb_adjoint_objective": [C.c_void_p],
        "tclb_adjoint_get_field": [C.c_void_p, C.c_void_p, C.c_char_p, C.c_void_p],
        "tclb_adjoint_want_setting": [C.c_void_p, C.c_void_p, C.c_char_p],
        "tclb_adjoint_setting_gradient": [C.c_void_p, C.c_void_p, C.c_char_p, C.POINTER(C.c_double)],
        "tclb_adjoint_set_field": [C.c_void_p, C.c_void_p, C.c_char_p, C.c_void_p],
    }.items():
        getattr(lib, fn).argtypes = args
    _core = lib
    return lib


def _check(rc: int):
    if rc != 0:
        raise NativeError(core().tclb_last_error().decode())


def _s(b) -> str:
    return b.decode() if b else ""


class Model:
    """A loaded model library and its metadata tables."""

    def __init__(self, variant: str, precision: str = "float", path: Optional[str] = None):
        lib = core()
        if path is None:
            p = LIBDIR / f"libtclb_{variant}_{precision}.so"
            if not p.exists():
                from . import build
                build.build_model_lib(variant, precision)
            path = str(p)
        self.path = path
        self.h = lib.tclb_model_load(path.encode())
        if not self.h:
            raise NativeError(lib.tclb_last_error().decode())
        self.d = lib.tclb_model_desc(self.h).contents
        d = self.d
        self.name = _s(d.name)
        self.precision = "double" if d.real_size == 8 else "float"
        self.real_dtype = np.float64 if d.real_size == 8 else np.float32
        self.flag_dtype = np.uint16 if d.flag_size == 2 else np.uint32
        self.fields = [_s(d.fields[i].name) for i in range(d.n_fields)]
        self.densities = [_s(d.densities[i].name) for i in range(d.n_densities)]
        self.settings = [_s(d.settings[i].name) for i in range(d.n_settings)]
        self.zonesettings = [_s(d.zonesettings[i].name) for i in range(d.n_zonesettings)]
        self.globals = [_s(d.globals[i].name) for i in range(d.n_globals)]
        self.quantities = {_s(d.quantities[i].name): bool(d.quantities[i].vector) for i in range(d.n_quantities)}
        self.quantity_units = {_s(d.quantities[i].name): _s(d.quantities[i].unit) for i in range(d.n_quantities)}
        self.nodetypes = {_s(d.nodetypes[i].name): (int(d.nodetypes[i].value), int(d.nodetypes[i].mask))
                          for i in range(d.n_nodetypes)}
        self.nodegroups = {_s(d.nodegroups[i].name): (int(d.nodegroups[i].mask), int(d.nodegroups[i].shift))
                           for i in range(d.n_nodegroups)}
        self.actions = {_s(d.actions[i].name): [int(d.actions[i].stages[k]) for k in range(d.actions[i].n_stages)]
                        for i in range(d.n_actions)}
        self.stages = [_s(d.stages[i].name) for i in range(d.n_stages)]
        self.options = {_s(d.options[i].name): int(d.options[i].value) for i in range(d.n_options)}
        self.setting_defaults = {_s(d.settings[i].name): _s(d.settings[i].default_value) for i in range(d.n_settings)}
        self.setting_defaults.update({_s(d.zonesettings[i].name): _s(d.zonesettings[i].default_value)
                                      for i in range(d.n_zonesettings)})
        self.objectives = [_s(d.objectives[i].name) for i in range(d.n_objectives)]
        self.zone_shift, self.zone_max = int(d.zone_shift), int(d.zone_max)
        self.storage = _s(d.storage_name)
        self.field_shift = [float(d.field_shift[i]) if d.storage_shifted else 0.0 for i in range(d.n_fields)]
        self.gy, self.gz = int(d.gy), int(d.gz)

    @property
    def bytes_per_update(self) -> int:
        return 2 * len(self.fields) * self.d.storage_size + self.d.flag_size

    def action_traffic(self, action: str = "Iteration") -> Dict[str, object]:
        """Compulsory memory traffic of one run of an action, from the stage tables the generator emitted (not an estimate):
        per stage, every field it loads as a density or reads through a stencil accessor counts once (its shifted copies
        come from the same rows), every field it saves counts once, plus the flag.  The reference's own accounting
        (src/main.cpp:127) assumes one stage that reads and writes every field: 2*nFields*sizeof + flag."""
        d = self.d
        dens_field = [int(d.densities[i].field_index) for i in range(d.n_densities)]
        per_stage, total = [], 0
        for s in self.actions[action]:
            st = d.stages[s]
            rd = set(dens_field[i] for i in range(d.n_densities) if st.load_density[i])
            rd |= set(f for f in range(d.n_fields) if st.read_field[f])
            wr = set(f for f in range(d.n_fields) if st.save_field[f])
            b = (len(rd) + len(wr)) * int(d.storage_size) + int(d.flag_size)
            per_stage.append({"stage": _s(st.name), "fields_read": len(rd), "fields_written": len(wr), "bytes": b})
            total += b
        return {"action": action, "bytes_per_update": total, "stages": per_stage}

    def objective(self, name: str, globals_values) -> tuple:
        """(value, gradient with respect to every Global) of a model objective at the given values of the Globals"""
        k = self.objectives.index(name)
        n = len(self.globals)
        g = (C.c_double * n)(*[float(globals_values[x]) if isinstance(globals_values, dict) else float(globals_values[i]) for i, x in enumerate(self.globals)])
        val, grad = (C.c_double * 1)(), (C.c_double * n)()
        self.d.objectives[k].fun(g, val, grad)
        return float(val[0]), np.array(list(grad))

    def flag(self, *names: str) -> int:
        v = 0
        for n in names:
            v |= self.nodetypes[n][0]
        return v


class Domain:
    """All subdomains owned by this process (see csrc/core/lattice.h)."""

    def __init__(self, model: Model, shape, backend: str = "cpu", nranks: int = 1,
                 local_ranks: Optional[Sequence[int]] = None, devices: Optional[Sequence[int]] = None,
                 apply_defaults: bool = True):
        lib = core()
        self.lib = lib
        self.model = model
        nx, ny, nz = (list(shape) + [1, 1])[:3]
        self.backend = BACKEND_CUDA if backend in ("cuda", "gpu", BACKEND_CUDA) else BACKEND_CPU
        lr = list(range(nranks)) if local_ranks is None else list(local_ranks)
        dv = list(devices) if devices is not None else [0]
        a_lr = (C.c_int * len(lr))(*lr)
        a_dv = (C.c_int * len(dv))(*dv)
        self.h = lib.tclb_domain_create(model.h, self.backend, nx, ny, nz, nranks, a_lr, len(lr), a_dv, len(dv))
        if not self.h:
            raise NativeError(lib.tclb_last_error().decode())
        self.nranks, self.local = nranks, lr
        divy, divz = C.c_int(), C.c_int()
        boxes = (C.c_int * (6 * nranks))()
        _check(lib.tclb_domain_decomposition(self.h, C.byref(divy), C.byref(divz), boxes))
        self.divy, self.divz = divy.value, divz.value
        self.boxes = [tuple(boxes[6 * r:6 * r + 6]) for r in range(nranks)]
        self.shape = (nx, ny, nz)
        if apply_defaults:
            for n, v in model.setting_defaults.items():
                try:
                    self.set_setting(n, float(v))
                except ValueError:
                    from .utils.units import UnitEnv
                    self.set_setting(n, UnitEnv().alt(v))

    @classmethod
    def from_handle(cls, model: Model, handle, nranks: int, local_ranks: Sequence[int]) -> "Domain":
        """wrap a lattice owned by a native Solver (borrowed handle: never destroyed from here)"""
        self = cls.__new__(cls)
        self.lib, self.model, self.h, self._borrowed = core(), model, handle, True
        self.backend = None
        self.nranks, self.local = nranks, list(local_ranks)
        divy, divz = C.c_int(), C.c_int()
        boxes = (C.c_int * (6 * nranks))()
        _check(self.lib.tclb_domain_decomposition(self.h, C.byref(divy), C.byref(divz), boxes))
        self.divy, self.divz = divy.value, divz.value
        self.boxes = [tuple(boxes[6 * r:6 * r + 6]) for r in range(nranks)]
        self.shape = (max(b[0] + b[3] for b in self.boxes), max(b[1] + b[4] for b in self.boxes), max(b[2] + b[5] for b in self.boxes))
        return self

    def close(self):
        if getattr(self, "h", None):
            if not getattr(self, "_borrowed", False):
                self.lib.tclb_domain_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- settings ----
    def set_setting(self, name: str, value: float, zone: Optional[int] = None):
        if zone is None:
            _check(self.lib.tclb_domain_set_setting(self.h, name.encode(), float(value)))
        else:
            _check(self.lib.tclb_domain_set_zone_setting(self.h, name.encode(), int(zone), float(value)))

    def get_setting(self, name: str) -> float:
        return float(self.lib.tclb_domain_get_setting(self.h, name.encode()))

    def set_zone_series(self, name: str, zone: int, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        _check(self.lib.tclb_domain_set_zone_series(self.h, name.encode(), zone, v.ctypes.data_as(C.POINTER(C.c_double)), len(v)))

    def set_halo_mode(self, direct: bool):
        _check(self.lib.tclb_domain_set_halo_mode(self.h, 1 if direct else 0))

    def use_graphs(self, on: bool = True):
        """run launch-only stretches of iterate() as a replayed CUDA graph (default on for CUDA domains)"""
        _check(self.lib.tclb_domain_use_graphs(self.h, 1 if on else 0))

    @property
    def graph_iterations(self) -> int:
        self.lib.tclb_domain_graph_iterations.restype = C.c_longlong
        return int(self.lib.tclb_domain_graph_iterations(self.h))

    def set_block(self, bx: int, by: int = 0):
        _check(self.lib.tclb_domain_set_block(self.h, bx, by))

    # ---- flags ----
    def set_flags(self, flags_global: np.ndarray):
        """flags_global: uint array [nz][ny][nx] of the whole lattice; scattered to local ranks."""
        nx, ny, nz = self.shape
        fl = np.ascontiguousarray(flags_global, dtype=np.uint32).reshape(nz, ny, nx)
        for r in self.local:
            x0, y0, z0, bx, by, bz = self.boxes[r]
            part = np.ascontiguousarray(fl[z0:z0 + bz, y0:y0 + by, x0:x0 + bx])
            _check(self.lib.tclb_domain_set_flags(self.h, r, part.ctypes.data))

    def get_flags(self) -> np.ndarray:
        nx, ny, nz = self.shape
        out = np.zeros((nz, ny, nx), dtype=np.uint32)
        for r in self.local:
            x0, y0, z0, bx, by, bz = self.boxes[r]
            part = np.zeros((bz, by, bx), dtype=np.uint32)
            _check(self.lib.tclb_domain_get_flags(self.h, r, part.ctypes.data))
            out[z0:z0 + bz, y0:y0 + by, x0:x0 + bx] = part
        return out

    # ---- execution ----
    def init(self):
        _check(self.lib.tclb_domain_init(self.h))

    def iterate(self, n: int, globals_on_last: bool = False, globals_every: bool = False):
        _check(self.lib.tclb_domain_iterate(self.h, int(n), int(globals_on_last), int(globals_every)))

    def run_action(self, name: str, globals_mode: int = GLOBALS_NONE):
        _check(self.lib.tclb_domain_run_action(self.h, name.encode(), globals_mode))

    def sync(self):
        _check(self.lib.tclb_domain_sync(self.h))

    @property
    def iter(self) -> int:
        return int(self.lib.tclb_domain_iter(self.h))

    def globals(self) -> Dict[str, float]:
        out = (C.c_double * max(1, len(self.model.globals)))()
        _check(self.lib.tclb_domain_get_globals(self.h, out))
        return {n: out[i] for i, n in enumerate(self.model.globals)}

    # ---- data ----
    def get_field(self, name: str) -> np.ndarray:
        nx, ny, nz = self.shape
        out = np.zeros((nz, ny, nx), dtype=self.model.real_dtype)
        for r in self.local:
            x0, y0, z0, bx, by, bz = self.boxes[r]
            part = np.zeros((bz, by, bx), dtype=self.model.real_dtype)
            _check(self.lib.tclb_domain_get_field(self.h, r, name.encode(), part.ctypes.data))
            out[z0:z0 + bz, y0:y0 + by, x0:x0 + bx] = part
        return out

    def set_field(self, name: str, values: np.ndarray):
        nx, ny, nz = self.shape
        v = np.ascontiguousarray(values, dtype=self.model.real_dtype).reshape(nz, ny, nx)
        for r in self.local:
            x0, y0, z0, bx, by, bz = self.boxes[r]
            part = np.ascontiguousarray(v[z0:z0 + bz, y0:y0 + by, x0:x0 + bx])
            _check(self.lib.tclb_domain_set_field(self.h, r, name.encode(), part.ctypes.data))

    def get_quantity(self, name: str, scale: float = 1.0) -> np.ndarray:
        nx, ny, nz = self.shape
        nc = 3 if self.model.quantities[name] else 1
        out = np.zeros((nz, ny, nx, nc), dtype=self.model.real_dtype)
        for r in self.local:
            x0, y0, z0, bx, by, bz = self.boxes[r]
            part = np.zeros((bz, by, bx, nc), dtype=self.model.real_dtype)
            _check(self.lib.tclb_domain_get_quantity(self.h, r, name.encode(), None, part.ctypes.data, float(scale)))
            out[z0:z0 + bz, y0:y0 + by, x0:x0 + bx] = part
        return out if nc == 3 else out[..., 0]

    def color(self, zplane: int = 0, scale: float = 1.0, rank: int = 0) -> np.ndarray:
        x0, y0, z0, bx, by, bz = self.boxes[rank]
        out = np.zeros((by, bx, 4), dtype=np.uint8)
        _check(self.lib.tclb_domain_color(self.h, rank, zplane, float(scale), out.ctypes.data))
        return out

    def pri_blocks(self, rank: int = 0):
        """elements of the 27 margin blocks of rank's checkpoint image (block1 .. block27 of the reference)"""
        out = (C.c_longlong * 27)()
        _check(self.lib.tclb_domain_pri_blocks(self.h, rank, out))
        return list(out)

    def save_pri(self, rank: int, path: str):
        _check(self.lib.tclb_domain_save_pri(self.h, rank, str(path).encode()))

    def load_pri(self, rank: int, path: str):
        _check(self.lib.tclb_domain_load_pri(self.h, rank, str(path).encode()))

    def snapshot_info(self, rank: int, snap: int = -1):
        shape = (C.c_int64 * 4)()
        gy, gz, cur = C.c_int(), C.c_int(), C.c_int()
        p = self.lib.tclb_domain_snapshot_ptr(self.h, rank, snap, shape, C.byref(gy), C.byref(gz), C.byref(cur))
        if not p:
            raise NativeError(self.lib.tclb_last_error().decode())
        return int(p), tuple(shape), gy.value, gz.value, cur.value

    def stream(self, rank: int) -> int:
        return int(self.lib.tclb_domain_stream(self.h, rank) or 0)


class Adjoint:
    """Discrete adjoint of a Domain (csrc/core/adjoint.h; the reference's <Adjoint> machinery).

    Needs the adjoint companion library (``tclb_b200.build.build_adjoint_lib(model, precision)``).  Typical use::

        adj = Adjoint(domain)
        adj.record(100)            # 100 primal iterations on tape, adj.objective = sum of the Objective global
        adj.reverse()              # backward sweep
        g = adj.get_field("w")     # dObjective/dw for a parameter field w, or dObjective/d(initial f) for a density

    One subdomain, Iteration actions of one stage, plain float/double storage."""

    def __init__(self, domain: "Domain"):
        self.domain = domain
        self.lib = domain.lib
        self.h = self.lib.tclb_adjoint_create(domain.h)
        if not self.h:
            raise NativeError(_s(self.lib.tclb_last_error()))

    def close(self):
        if self.h:
            self.lib.tclb_adjoint_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def record(self, n: int):
        _check(self.lib.tclb_adjoint_record(self.h, int(n)))

    def reverse(self):
        _check(self.lib.tclb_adjoint_reverse(self.h))

    def steady(self, n: int):
        _check(self.lib.tclb_adjoint_steady(self.h, int(n)))

    def zero(self):
        _check(self.lib.tclb_adjoint_zero(self.h))

    def set_tape_frames(self, n: int):
        """keep at most n stored states: checkpoints at a doubling stride during record(), recomputation in reverse()"""
        _check(self.lib.tclb_adjoint_set_tape_frames(self.h, int(n)))

    def tape_frames(self) -> int:
        return int(self.lib.tclb_adjoint_tape_frames(self.h))

    @property
    def objective(self) -> float:
        return float(self.lib.tclb_adjoint_objective(self.h))

    def want_setting_gradients(self, names):
        """ask for dObjective/d(setting) of these non-zonal settings (before record())"""
        _check(self.lib.tclb_adjoint_want_setting(self.h, self.domain.h, ",".join(names).encode()))

    def setting_gradient(self, name: str) -> float:
        out = C.c_double(0.0)
        _check(self.lib.tclb_adjoint_setting_gradient(self.h, self.domain.h, name.encode(), C.byref(out)))
        return out.value

    def get_field(self, name: str) -> np.ndarray:
        nx, ny, nz = self.domain.shape
        out = np.zeros((nz, ny, nx), dtype=self.domain.model.real_dtype)
        _check(self.lib.tclb_adjoint_get_field(self.h, self.domain.h, name.encode(), out.ctypes.data))
        return out

    def set_field(self, name: str, values: np.ndarray):
        nx, ny, nz = self.domain.shape
        v = np.ascontiguousarray(values, dtype=self.domain.model.real_dtype).reshape(nz, ny, nx)
        _check(self.lib.tclb_adjoint_set_field(self.h, self.domain.h, name.encode(), v.ctypes.data))
