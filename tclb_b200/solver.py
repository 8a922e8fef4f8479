"""Case runner with in-process Python coupling.

``Solver(model).load("case.xml").run()`` executes a TCLB ``<CLBConfig>`` case through the native handler tree
(csrc/handlers) exactly like ``bin/tclb``; in addition ``<RunPython>`` elements run here, in this interpreter, with an
object ``Solver`` that offers what the reference's embedded R/Python gets (src/Handlers/cbRunR.cpp:484-520, 751-845):

    Solver.Settings["nu"] = 0.01          Solver.Settings["Velocity", "Inlet"] = 0.02     (zone by name)
    Solver.Fields["BC[0]"] = array        a = Solver.Fields["f[3]"]                       ([nz][ny][nx] numpy arrays)
    Solver.Quantities["U"]                Solver.Globals["Flux"]
    Solver.Geometry.X / .Y / .Z           cell-centred coordinates
    Solver.Geometry["BOUNDARY"][sel] = "Wall"      node-type groups by name; "None" clears the group
    Solver.Actions.Iterate(100)           Solver.Actions.Run("InitFields")
    Solver.Info.Iteration                 Solver.Info.Model

``<RunR>`` elements are handed to the same hook and run in this repository's R-subset interpreter (`Solver._run_r`).
"""
from __future__ import annotations

import ctypes as C
import traceback
from typing import Dict, List, Optional, Sequence

import numpy as np

from .runtime import BACKEND_CPU, BACKEND_CUDA, Domain, Model, NativeError, _check, core

_SCRIPT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_int)


class _Settings:
    def __init__(self, s: "Solver"):
        self._s = s

    def __getitem__(self, key):
        name = key[0] if isinstance(key, tuple) else key
        return self._s.domain.get_setting(name)

    def __setitem__(self, key, value):
        d = self._s.domain
        if isinstance(key, tuple):
            name, zone = key
            z = self._s.zones().get(zone, zone) if isinstance(zone, str) else int(zone)
            d.set_setting(name, float(value), zone=int(z))
        else:
            d.set_setting(key, float(value))

    def keys(self):
        m = self._s.model
        return list(m.settings) + list(m.zonesettings)


class _Fields:
    def __init__(self, s: "Solver"):
        self._s = s

    def __getitem__(self, name):
        return self._s.domain.get_field(name)

    def __setitem__(self, name, value):
        nx, ny, nz = self._s.domain.shape
        self._s.domain.set_field(name, np.broadcast_to(np.asarray(value, dtype=np.float64), (nz, ny, nx)))

    def keys(self):
        return list(self._s.model.fields)


class _Quantities:
    def __init__(self, s: "Solver"):
        self._s = s

    def __getitem__(self, name):
        return self._s.domain.get_quantity(name)

    def keys(self):
        return list(self._s.model.quantities)


class _Globals:
    def __init__(self, s: "Solver"):
        self._s = s

    def __getitem__(self, name):
        return self._s.domain.globals()[name]

    def keys(self):
        return list(self._s.model.globals)


class _GroupView:
    """one node-type group of the flag array, read and written by node-type NAME"""

    def __init__(self, geo: "_Geometry", group: str):
        self._g, self._group = geo, group

    def _mask(self):
        return self._g._s.model.nodegroups[self._group][0]

    def names(self) -> np.ndarray:
        m = self._g._s.model
        fl = self._g.flags()
        mask = self._mask()
        out = np.full(fl.shape, "None", dtype=object)
        for name, (val, msk) in m.nodetypes.items():
            if msk == mask:
                out[(fl & mask) == val] = name
        return out

    def __getitem__(self, sel):
        return self.names()[sel]

    def __setitem__(self, sel, name):
        m = self._g._s.model
        fl = self._g.flags()
        mask = self._mask()
        val = 0 if name in ("None", None) else m.nodetypes[name][0]
        fl[sel] = (fl[sel] & ~np.uint32(mask)) | np.uint32(val)
        self._g._s.domain.set_flags(fl)


class _Geometry:
    def __init__(self, s: "Solver"):
        self._s = s

    def flags(self) -> np.ndarray:
        return self._s.domain.get_flags()

    def _coord(self, axis):
        nx, ny, nz = self._s.domain.shape
        z, y, x = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
        return (x, y, z)[axis] + 0.5

    X = property(lambda self: self._coord(0))
    Y = property(lambda self: self._coord(1))
    Z = property(lambda self: self._coord(2))

    def __getitem__(self, group):
        if group == "SETTINGZONE":
            return _ZoneView(self)
        return _GroupView(self, group)

    def __getattr__(self, group):
        if group.startswith("_"):
            raise AttributeError(group)
        return self[group]


class _ZoneView:
    def __init__(self, geo: _Geometry):
        self._g = geo

    def __setitem__(self, sel, zone):
        s = self._g._s
        z = s.zones().get(zone, zone) if isinstance(zone, str) else int(zone)
        fl = self._g.flags()
        shift = s.model.zone_shift
        fl[sel] = (fl[sel] & np.uint32((1 << shift) - 1)) | np.uint32(int(z) << shift)
        s.domain.set_flags(fl)

    def __getitem__(self, sel):
        return (self._g.flags() >> self._g._s.model.zone_shift)[sel]


class _Actions:
    def __init__(self, s: "Solver"):
        self._s = s

    def Iterate(self, n: int = 1, globals_every: bool = False):
        self._s.domain.iterate(int(n), globals_on_last=True, globals_every=globals_every)

    def Run(self, name: str):
        self._s.domain.run_action(name)

    def Init(self):
        self._s.domain.init()


class _Info:
    def __init__(self, s: "Solver"):
        self._s = s

    Iteration = property(lambda self: core().tclb_solver_iter(self._s.h))
    Model = property(lambda self: self._s.model.name)
    OutputPrefix = property(lambda self: core().tclb_solver_outpath(self._s.h).decode())


def _dedent(code: str) -> str:
    """common indentation removed; an XML parser may already have trimmed the first line, so it is treated separately"""
    import textwrap
    lines = code.expandtabs(4).split("\n")
    while lines and not lines[0].strip():
        lines.pop(0)
    if not lines:
        return ""
    first, rest = lines[0], lines[1:]
    body = textwrap.dedent("\n".join(rest))
    if first[:1].strip():                       # first line starts in column 0: dedent the rest on its own
        return first + "\n" + body
    return textwrap.dedent("\n".join(lines))


class _ZoneNames:
    def __init__(self, s: "Solver"):
        self._s = s

    def get(self, name, default=None):
        z = self._s.lib.tclb_solver_zone_id(self._s.h, str(name).encode())
        return z if z >= 0 else default

    def __getitem__(self, name):
        z = self.get(name)
        if z is None:
            raise KeyError(name)
        return z


class Solver:
    def __init__(self, model: str, precision: str = "double", backend: str = "cpu", nranks: int = 1,
                 local_ranks: Optional[Sequence[int]] = None, devices: Optional[Sequence[int]] = None, direct_halo: bool = True):
        lib = core()
        lib.tclb_solver_create.restype = C.c_void_p
        lib.tclb_solver_create.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int]
        lib.tclb_solver_destroy.argtypes = [C.c_void_p]
        lib.tclb_solver_load_file.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_char_p), C.c_int]
        lib.tclb_solver_load_string.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p), C.c_int]
        lib.tclb_solver_run.argtypes = [C.c_void_p]
        lib.tclb_solver_domain.restype = C.c_void_p
        lib.tclb_solver_domain.argtypes = [C.c_void_p]
        lib.tclb_solver_iter.argtypes = [C.c_void_p]
        lib.tclb_solver_outpath.restype = C.c_char_p
        lib.tclb_solver_outpath.argtypes = [C.c_void_p]
        lib.tclb_solver_set_script_hook.argtypes = [C.c_void_p, _SCRIPT_FN, C.c_void_p]
        lib.tclb_solver_zone_id.argtypes = [C.c_void_p, C.c_char_p]
        self.lib = lib
        self.model = Model(model, precision)
        be = BACKEND_CUDA if backend in ("cuda", "gpu") else BACKEND_CPU
        lr = list(range(nranks)) if local_ranks is None else list(local_ranks)
        dv = list(devices) if devices is not None else [0]
        self.h = lib.tclb_solver_create(model.encode(), precision.encode(), be, nranks, (C.c_int * len(lr))(*lr), len(lr),
                                        (C.c_int * len(dv))(*dv), len(dv), 1 if direct_halo else 0)
        if not self.h:
            raise NativeError(lib.tclb_last_error().decode())
        self.nranks, self.local = nranks, lr
        self.domain: Optional[Domain] = None
        self.namespace: Dict[str, object] = {}
        self.error: Optional[str] = None
        self.Settings, self.Fields, self.Quantities = _Settings(self), _Fields(self), _Quantities(self)
        self.Globals, self.Geometry, self.Actions, self.Info = _Globals(self), _Geometry(self), _Actions(self), _Info(self)
        self._hook = _SCRIPT_FN(self._on_script)
        _check(lib.tclb_solver_set_script_hook(self.h, self._hook, None))

    # ---- set-up ----
    def load(self, case_file: str, xpath_args: Sequence[str] = ()):
        arr = (C.c_char_p * max(1, len(xpath_args)))(*[a.encode() for a in xpath_args])
        _check(self.lib.tclb_solver_load_file(self.h, case_file.encode(), arr, len(xpath_args)))
        self._bind_domain()
        return self

    def load_string(self, xml: str, name: str = "case", xpath_args: Sequence[str] = ()):
        arr = (C.c_char_p * max(1, len(xpath_args)))(*[a.encode() for a in xpath_args])
        _check(self.lib.tclb_solver_load_string(self.h, xml.encode(), name.encode(), arr, len(xpath_args)))
        self._bind_domain()
        return self

    def _bind_domain(self):
        dh = self.lib.tclb_solver_domain(self.h)
        if not dh:
            raise NativeError("the case did not create a lattice")
        self.domain = Domain.from_handle(self.model, dh, self.nranks, self.local)

    def zones(self) -> Dict[str, int]:
        """setting-zone names of the case (DefaultZone = 0, then in order of first appearance, like the reference)"""
        return _ZoneNames(self)

    def run(self):
        rc = self.lib.tclb_solver_run(self.h)
        if rc != 0:
            raise NativeError((self.error + "\n" if self.error else "") + self.lib.tclb_last_error().decode())
        return self

    def close(self):
        if getattr(self, "h", None):
            if self.domain is not None:
                self.domain.h = None            # borrowed handle: the solver owns the lattice
            self.lib.tclb_solver_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- <RunR> ----
    def _run_r(self, src: str, iteration: int):
        """<RunR> without R: the block runs in this repository's R-subset interpreter (tclb_b200/gen/rinterp.py, the one that
        executes the models' Dynamics.R) with a `Solver` list of the same shape the reference's RInside bridge offers
        (src/Handlers/cbRunR.cpp:484-520, 751-845): Solver$Settings$<name>, Solver$Globals$<name>, Solver$Quantities$<name>,
        Solver$Fields$<name>, Solver$Geometry$X|Y|Z, Solver$Geometry$<GROUP> (node-type names per node), Solver$Actions$<name>(), Solver$Iteration.  Lattice arrays are vectors in the lattice's memory
        order (x fastest; a vector quantity is 3 x nodes).  Only what the block mentions is materialised; after the block,
        fields, settings and node types it changed are written back.  The language coverage is the interpreter's (vector arithmetic,
        data frames, apply-family, control flow, write.csv), not all of R."""
        import re

        from .gen.rinterp import Interp
        from .gen.rlang import RList, vec
        m, d = self.model, self.domain
        used = lambda kind: sorted(set(a or b for a, b in re.findall(r"Solver\$" + kind + r"\$(?:`([^`]+)`|([A-Za-z_.][\w.]*))", src)))
        S = RList()
        st = RList()
        for k in list(m.settings) + list(m.zonesettings):
            st[k] = vec(float(d.get_setting(k)))
        S["Settings"] = st
        gl = RList()
        if "Solver$Globals" in src:
            for k, v in d.globals().items():
                gl[k] = vec(float(v))
        S["Globals"] = gl
        q = RList()
        for k in used("Quantities"):
            if k not in m.quantities:
                raise RuntimeError(f"<RunR>: the model has no quantity {k}")
            a = d.get_quantity(k)
            q[k] = np.ascontiguousarray(a, dtype=np.float64).reshape(-1, 3).T.copy() if m.quantities[k] else np.asarray(a, dtype=np.float64).ravel()
        S["Quantities"] = q
        f = RList()
        for k in used("Fields"):
            if k not in m.fields:
                raise RuntimeError(f"<RunR>: the model has no field {k}")
            f[k] = np.asarray(d.get_field(k), dtype=np.float64).ravel()
        S["Fields"] = f
        g = RList()
        geo = _Geometry(self)
        for k in used("Geometry"):
            if k in ("X", "Y", "Z"):
                g[k] = np.asarray(getattr(geo, k), dtype=np.float64).ravel()
            elif k in m.nodegroups and k != "SETTINGZONE":
                g[k] = np.array([str(x) for x in _GroupView(geo, k).names().ravel()], dtype=object)    # node-type names per node
            else:
                raise RuntimeError(f"<RunR>: Solver$Geometry${k} is not available here (X, Y, Z and the node-type groups are)")
        S["Geometry"] = g
        S["Iteration"] = vec(float(iteration))
        before_groups = {k: np.array(v, copy=True) for k, v in g.items() if k not in ("X", "Y", "Z")}
        before_fields = {k: np.array(v, copy=True) for k, v in f.items()}
        before_settings = {k: float(np.asarray(v).ravel()[0]) for k, v in st.items()}
        I = getattr(self, "_r_interp", None)
        if I is None:
            I = self._r_interp = Interp()
        nx, ny, nz = d.shape

        def flush():
            """what the R code changed so far goes to the lattice (end of the block, and before an action it calls)"""
            S2 = I.globals.get("Solver")
            for k, v in S2["Fields"].items():
                a = np.asarray(v, dtype=np.float64).ravel()
                if k not in m.fields:
                    raise RuntimeError(f"<RunR>: the model has no field {k}")
                if k not in before_fields or a.shape != before_fields[k].shape or not np.array_equal(a, before_fields[k]):
                    d.set_field(k, np.resize(a, nx * ny * nz).reshape(nz, ny, nx))
                    before_fields[k] = np.resize(a, nx * ny * nz).copy()
            for k, old_names in before_groups.items():
                now = np.array([str(x) for x in np.asarray(S2["Geometry"][k], dtype=object).ravel()], dtype=object)
                if now.shape != old_names.shape:
                    raise RuntimeError(f"<RunR>: Solver$Geometry${k} changed its length")
                changed = now != old_names
                if changed.any():
                    fl = geo.flags()
                    flat = fl.reshape(-1)
                    mask = np.uint32(m.nodegroups[k][0])
                    for name in sorted(set(now[changed].tolist())):
                        if name not in ("None", "") and (name not in m.nodetypes or m.nodetypes[name][1] != int(mask)):
                            raise RuntimeError(f"<RunR>: {name} is not a node type of group {k}")
                        val = np.uint32(0 if name in ("None", "") else m.nodetypes[name][0])
                        sel = changed & (now == name)
                        flat[sel] = (flat[sel] & ~mask) | val
                    d.set_flags(fl)
                    before_groups[k] = now
            for k, v in S2["Settings"].items():
                x = float(np.asarray(v).ravel()[0])
                if k not in before_settings:
                    raise RuntimeError(f"<RunR>: the model has no setting {k}")
                if x != before_settings[k] and not (x != x and before_settings[k] != before_settings[k]):     # NaN stays NaN: not a change
                    d.set_setting(k, x)
                    before_settings[k] = x

        def refresh():
            """after an action: the fields and quantities the block holds show the new state"""
            S2 = I.globals.get("Solver")
            for k in list(S2["Fields"].keys()):
                S2["Fields"][k] = np.asarray(d.get_field(k), dtype=np.float64).ravel()
                before_fields[k] = np.array(S2["Fields"][k], copy=True)
            for k in list(S2["Quantities"].keys()):
                a = d.get_quantity(k)
                S2["Quantities"][k] = np.ascontiguousarray(a, dtype=np.float64).reshape(-1, 3).T.copy() if m.quantities[k] else np.asarray(a, dtype=np.float64).ravel()

        # Solver$Actions$<name>(): Init, Iterate(n) and the model's actions (e.g. InitFromFields), applied to what the block wrote so far
        acts = RList()

        def make_action(name):
            def run(I_, a, k):
                flush()
                if name == "Iterate":
                    d.iterate(int(np.asarray(a[0]).ravel()[0]) if a else 1)
                elif name == "Init":
                    d.init()
                else:
                    d.run_action(name)
                refresh()
                from .gen.rlang import NULL
                return NULL
            return run
        for name in ["Iterate", "Init"] + [a_ for a_ in m.actions if a_ not in ("Iteration", "Init")]:
            acts[name] = make_action(name)
        S["Actions"] = acts
        I.globals.set("Solver", S)
        I.run(src, I.globals)
        flush()

    # ---- <RunPython> ----
    def _on_script(self, user, element, code, iteration) -> int:
        try:
            if element.decode() == "RunR":
                self._run_r(_dedent(code.decode()), iteration)
                return 0
            src = _dedent(code.decode())
            ns = self.namespace
            ns.setdefault("np", np)
            ns["Solver"] = self
            exec(compile(src, f"<RunPython @ iteration {iteration}>", "exec"), ns)
            return 0
        except Exception:
            self.error = traceback.format_exc()
            print(self.error)
            return 1
