"""Model description language + derived tables.

This is the counterpart of the reference's model DSL and table builder
(``/root/reference/src/conf.R``: AddDensity/AddField/AddSetting/AddGlobal/AddQuantity/
AddNodeType/AddStage/AddAction :54-360, node-type bit packing :600-698, Globals ordering
:740-751, automatic ``<Global>InObj`` zonal settings :753-761, ``Threshold`` :769, default
stages/actions :459-472).  The call names and keyword arguments are kept identical so that
a ``Dynamics.R`` file evaluated by :mod:`tclb_b200.gen.rlang` maps 1:1 onto these methods,
but the implementation (plain Python dataclasses, no data frames) is new.
"""
from __future__ import annotations

import math
import re
from dataclasses import dataclass, field
from typing import Any, Dict, Iterable, List, Optional, Sequence, Union


def _aslist(x) -> list:
    if isinstance(x, (list, tuple)):
        return list(x)
    try:
        import numpy as np
        if isinstance(x, np.ndarray):
            return [v.item() if hasattr(v, "item") else v for v in x.tolist()] if x.ndim else [x.item()]
    except Exception:
        pass
    return [x]


def nicename(n: str) -> str:
    return re.sub(r"[\]\[ ]", "", n)


@dataclass
class Density:
    name: str
    field: str
    dx: int = 0
    dy: int = 0
    dz: int = 0
    comment: str = ""
    group: str = ""
    parameter: bool = False
    average: bool = False
    adjoint: bool = False
    default: Optional[float] = None
    load: Dict[str, bool] = field(default_factory=dict)      # stage -> loaded?
    sym: List[str] = field(default_factory=lambda: ["", "", ""])   # mirror partner per axis (autosym), "-name" = with sign


@dataclass
class Field:
    name: str
    minx: int = 0
    maxx: int = 0
    miny: int = 0
    maxy: int = 0
    minz: int = 0
    maxz: int = 0
    comment: str = ""
    group: str = ""
    parameter: bool = False
    average: bool = False
    adjoint: bool = False
    shift: Optional[float] = None
    optimise_for_static_access: bool = True
    non_mandatory: bool = False
    save: Dict[str, bool] = field(default_factory=dict)      # stage -> saved?
    read: Dict[str, Optional[bool]] = field(default_factory=dict)
    sym: List[str] = field(default_factory=lambda: ["", "", ""])   # mirror partner per axis (autosym), "-name" = with sign

    @property
    def simple_access(self) -> bool:
        return (self.maxx - self.minx + 1) * (self.maxy - self.miny + 1) * (self.maxz - self.minz + 1) == 1


@dataclass
class Setting:
    name: str
    comment: str = ""
    default: Any = 0
    unit: str = "1"
    adjoint: bool = False
    derived: Optional[str] = None
    equation: Optional[str] = None
    zonal: bool = False
    preload: bool = True


@dataclass
class Global:
    name: str
    var: str = ""
    comment: str = ""
    unit: str = "1"
    adjoint: bool = False
    op: str = "SUM"
    base: float = 0.0


@dataclass
class Quantity:
    name: str
    unit: str = "1"
    vector: bool = False
    comment: str = ""
    adjoint: bool = False


@dataclass
class NodeType:
    name: str
    group: str
    index: int = 0
    value: int = 0
    mask: int = 0
    shift: int = 0
    bits: int = 0
    capacity: int = 0
    save: bool = True


@dataclass
class NodeTypeGroup:
    name: str
    mask: int
    shift: int
    bits: int
    capacity: int
    save: bool


@dataclass
class Stage:
    name: str
    main: str
    load_densities: Any = False
    save_fields: Any = False
    read_fields: Any = None
    can_overwrite: bool = False
    fixed_point: bool = False
    particle: bool = False


@dataclass
class Action:
    name: str
    stages: List[str]


class ModelError(Exception):
    pass


class ModelDef:
    """Collects the declarative description of one model and derives every table the
    runtime needs.  ``options`` mirrors the reference's ``Options$<name>`` flags."""

    def __init__(self, name: str, options: Optional[Dict[str, Any]] = None, adjoint: int = 0):
        self.name = name
        self.options: Dict[str, Any] = dict(options or {})
        self.adjoint = adjoint
        self.densities: List[Density] = []
        self.fields: List[Field] = []
        self.settings: List[Setting] = []
        self.zonesettings: List[Setting] = []
        self.globals: List[Global] = []
        self.quantities: List[Quantity] = []
        self.nodetypes: List[NodeType] = []
        self.nodetypegroups: List[NodeTypeGroup] = []
        self.stages: List[Stage] = []
        self.actions: List[Action] = []
        self.objectives: Dict[str, Any] = {}
        self.description: Optional[Dict[str, str]] = None
        self.permissive_access = False
        self.part_margin: Optional[float] = None
        self.code: str = ""           # node physics ("Dynamics.c" text)
        self.finalized = False
        # derived
        self.flag_bits = 16
        self.zone_bits = 0
        self.zone_shift = 0
        self.zone_max = 0
        self.sum_globals = 0
        self.in_obj_offset = 0

    # ---- declaration API (names follow the reference DSL) ---------------------------------
    def SetOptions(self, **kw):
        for k, v in kw.items():
            k = k.replace(".", "_")
            if k == "permissive_access":
                self.permissive_access = bool(v)
            elif k == "PartMargin":
                self.part_margin = float(v)
            else:
                raise ModelError(f"Unknown option in SetOptions: {k}")

    def AddDensity(self, name, dx=0, dy=0, dz=0, comment="", field=None, adjoint=False, group=None,
                   parameter=False, average=False, default=None, sym=None, shift=None, **kw):
        names = _aslist(name)
        n = len(names)
        def rep(v):
            v = _aslist(v)
            return [v[i % len(v)] for i in range(n)]
        dxs, dys, dzs = rep(dx), rep(dy), rep(dz)
        comments = rep(comment)
        fields_ = rep(field) if field is not None else names
        shifts = rep(shift)
        if average and default is None:
            default = 0.0
        for i, nm in enumerate(names):
            grp = group if group is not None else nm
            if parameter and (dxs[i] != 0 and dys[i] != 0 and dzs[i] != 0):
                raise ModelError("Parameters cannot be streamed (AddDensity)")
            d = Density(name=nm, field=fields_[i], dx=int(dxs[i]), dy=int(dys[i]), dz=int(dzs[i]),
                        comment=comments[i] or nm, group=grp, parameter=bool(parameter),
                        average=bool(average), adjoint=bool(adjoint),
                        default=None if default is None else float(default))
            if sym is not None:
                sv = [str(x) for x in _aslist(sym)]
                if len(sv) != 3:
                    raise ModelError("sym provided to AddDensity have to be a vector of length 3")
                d.sym = sv
            self.densities.append(d)
            self.AddField(name=d.field, dx=-d.dx, dy=-d.dy, dz=-d.dz, comment=d.comment,
                          adjoint=adjoint, group=grp, parameter=parameter, average=average,
                          shift=shifts[i], sym=sym, **kw)

    def AddField(self, name, stencil2d=None, stencil3d=None, dx=0, dy=0, dz=0, comment="",
                 adjoint=False, group=None, parameter=False, average=False, sym=None, shift=None,
                 optimise_for_static_access=True, non_mandatory=False, **kw):
        if "non.mandatory" in kw:
            non_mandatory = kw.pop("non.mandatory")
        for nm in _aslist(name):
            def lo(v, *st):
                c = [int(x) for x in _aslist(v)] + [-int(s) for s in st if s is not None]
                return min(c)
            def hi(v, *st):
                c = [int(x) for x in _aslist(v)] + [int(s) for s in st if s is not None]
                return max(c)
            f = Field(name=nm,
                      minx=lo(dx, stencil2d, stencil3d), maxx=hi(dx, stencil2d, stencil3d),
                      miny=lo(dy, stencil2d, stencil3d), maxy=hi(dy, stencil2d, stencil3d),
                      minz=lo(dz, stencil3d), maxz=hi(dz, stencil3d),
                      comment=comment or nm, group=group if group is not None else nm,
                      parameter=bool(parameter), average=bool(average), adjoint=bool(adjoint),
                      shift=None if shift in (None, 0) else float(shift),
                      optimise_for_static_access=bool(optimise_for_static_access),
                      non_mandatory=bool(non_mandatory))
            if sym is not None:
                sv = [str(x) for x in _aslist(sym)]
                if len(sv) == 3:
                    f.sym = sv
            old = self.field(nm, missing_ok=True)
            if old is not None:
                old.minx = min(old.minx, f.minx); old.maxx = max(old.maxx, f.maxx)
                old.miny = min(old.miny, f.miny); old.maxy = max(old.maxy, f.maxy)
                old.minz = min(old.minz, f.minz); old.maxz = max(old.maxz, f.maxz)
            else:
                if self.stages:
                    raise ModelError("Field added after a Stage was declared")
                self.fields.append(f)

    def AddSetting(self, name, comment=None, default=0, unit="1", adjoint=False, derived=None,
                   equation=None, zonal=False, preload=True, **der):
        if unit == "":
            raise ModelError("Empty unit in AddSetting not allowed")
        if derived is None:
            if len(der) == 1:
                derived, equation = next(iter(der.items()))
                equation = str(equation)
            elif len(der) > 1:
                raise ModelError("Only one derived setting allowed in AddSetting")
        elif equation is None:
            raise ModelError("'derived' provided, but no 'equation' in AddSetting")
        s = Setting(name=name, comment=comment or name, default=default, unit=unit, adjoint=bool(adjoint),
                    derived=derived, equation=equation, zonal=bool(zonal), preload=bool(preload))
        (self.zonesettings if zonal else self.settings).append(s)

    def AddGlobal(self, name, var=None, comment="", unit="1", adjoint=False, op="SUM", base=0.0):
        if op not in ("SUM", "MAX"):
            raise ModelError("Operation (op) in AddGlobal has to be SUM or MAX")
        self.globals.append(Global(name=name, var=var or name, comment=comment or name, unit=unit,
                                   adjoint=bool(adjoint), op=op, base=float(base)))

    def AddQuantity(self, name, unit="1", vector=False, comment="", adjoint=False):
        self.quantities.append(Quantity(name=name, unit=unit, vector=bool(vector),
                                        comment=comment or name, adjoint=bool(adjoint)))

    def AddNodeType(self, name, group):
        if not any(n.name == name and n.group == group for n in self.nodetypes):
            self.nodetypes.append(NodeType(name=name, group=group))

    def AddDescription(self, short, long=None):
        self.description = {"name": self.name, "short": short, "long": long or short}

    def AddStage(self, name, main=None, load_densities=False, save_fields=False, read_fields=None,
                 can_overwrite=False, default=False, fixedPoint=False, particle=False,
                 particle_margin=None, **kw):
        # dotted R names
        load_densities = kw.pop("load.densities", load_densities)
        save_fields = kw.pop("save.fields", save_fields)
        read_fields = kw.pop("read.fields", read_fields)
        can_overwrite = kw.pop("can.overwrite", can_overwrite)
        particle_margin = kw.pop("particle.margin", particle_margin)
        if any(s.name == name for s in self.stages):
            if default:
                return
            raise ModelError("Two stages defined with the same name")
        st = Stage(name=name, main=main or name, load_densities=load_densities, save_fields=save_fields,
                   read_fields=read_fields, can_overwrite=bool(can_overwrite), fixed_point=bool(fixedPoint),
                   particle=bool(particle))
        self.stages.append(st)
        if particle_margin is not None:
            self.part_margin = max(self.part_margin or 0.0, float(particle_margin))

        def selection(tab, sel, what):
            names = [t.name for t in tab]
            if sel is None:
                return [None] * len(tab)
            if isinstance(sel, bool):
                return [sel] * len(tab)
            sel = _aslist(sel)
            if all(isinstance(s, str) for s in sel):
                for s in sel:
                    if s not in names:
                        raise ModelError(f"{what} name '{s}' not found in AddStage({name})")
                return [n in sel for n in names]
            if all(isinstance(s, (bool,)) for s in sel) and len(sel) == len(tab):
                return list(sel)
            raise ModelError(f"invalid {what} selection in AddStage({name})")

        for d, v in zip(self.densities, selection(self.densities, load_densities, "load")):
            d.load[name] = bool(v)
        for f, v in zip(self.fields, selection(self.fields, save_fields, "save")):
            f.save[name] = bool(v)
        for f, v in zip(self.fields, selection(self.fields, read_fields, "read")):
            f.read[name] = v

    def AddAction(self, name, stages):
        self.actions.append(Action(name=name, stages=[str(s) for s in _aslist(stages)]))

    def AddObjective(self, name, expr):
        self.objectives[name] = expr

    # ---- lookups -------------------------------------------------------------------------
    def field(self, name, missing_ok=False) -> Optional[Field]:
        for f in self.fields:
            if f.name == name:
                return f
        if missing_ok:
            return None
        raise KeyError(name)

    def field_index(self, name) -> int:
        for i, f in enumerate(self.fields):
            if f.name == name:
                return i
        raise KeyError(name)

    def stage(self, name) -> Stage:
        for s in self.stages:
            if s.name == name:
                return s
        raise KeyError(name)

    def density_dicts(self, group: Optional[str] = None) -> List[dict]:
        return [dict(name=d.name, dx=d.dx, dy=d.dy, dz=d.dz, group=d.group)
                for d in self.densities if group is None or d.group == group]

    def group_names(self, group: str) -> List[str]:
        return [d.name for d in self.densities if d.group == group]

    def group_velocities(self, group: str, dim: int = 3):
        import numpy as np
        return np.array([[d.dx, d.dy, d.dz][:dim] for d in self.densities if d.group == group], dtype=int)

    # ---- derivation ----------------------------------------------------------------------
    def finalize(self) -> "ModelDef":
        if self.finalized:
            return self
        if not self.fields:
            raise ModelError("The model has to have at least one Field/Density")
        for g in list(self.globals):
            self.objectives.setdefault(g.name, g.name)
        if self.description is None:
            self.AddDescription(self.name)
        self._autosym()
        # default actions / stages
        if not any(a.name == "Iteration" for a in self.actions):
            self.AddAction("Iteration", ["BaseIteration"])
        if not any(a.name == "Init" for a in self.actions):
            self.AddAction("Init", ["BaseInit"])
        allst = [s for a in self.actions for s in a.stages]
        if "BaseIteration" in allst and not any(s.name == "BaseIteration" for s in self.stages):
            self.AddStage(main="Run", name="BaseIteration", load_densities=True, save_fields=True, default=True)
        if "BaseInit" in allst and not any(s.name == "BaseInit" for s in self.stages):
            self.AddStage(main="Init", name="BaseInit", load_densities=False, save_fields=True, default=True)
        self._check_actions()
        for f in self.fields:
            for s in self.stages:
                if self.permissive_access or f.read.get(s.name) is None:
                    f.read[s.name] = True
        self._pack_nodetypes()
        if not self.adjoint:
            # the reference compiles adjoint quantities only under #ifdef ADJOINT (src/LatticeContainer.inc.cpp.Rt:429)
            self.quantities = [q for q in self.quantities if not q.adjoint]
        # adjoint auto settings / globals
        if self.adjoint == 1:
            for s in list(self.settings):
                self.AddGlobal(name=f"{s.name}_D", var=f"{s.name}b",
                               comment=f"Gradient of objective with respect to [{s.comment}]", adjoint=True)
            self.AddSetting(name="Descent", comment="Optimization Descent", adjoint=True)
            self.AddSetting(name="GradientSmooth", comment="Gradient smoothing in OptSolve", adjoint=True)
            self.AddGlobal(name="AdjointRes", comment="square L2 norm of adjoint change", adjoint=True)
        elif getattr(self, "adjoint_requested", 0) == 1:
            # primal build of an ADJOINT=1 model: its case files still set the optimisation step (<OptSolve> reads it)
            self.AddSetting(name="Descent", comment="Optimization Descent", adjoint=True)
            self.AddSetting(name="GradientSmooth", comment="Gradient smoothing in OptSolve", adjoint=True)
        self.AddGlobal(name="Objective", comment="Objective function")
        c = sorted([g for g in self.globals if g.op != "SUM"], key=lambda g: g.op)
        b = [g for g in self.globals if g.name == "Objective"]
        a = [g for g in self.globals if g.op == "SUM" and g.name != "Objective"]
        self.globals = a + b + c
        self.sum_globals = sum(1 for g in self.globals if g.op == "SUM")
        for g in a:
            if not g.adjoint:
                self.AddSetting(name=f"{g.name}InObj", comment=f"Weight of [{g.comment}] in objective",
                                preload=False, adjoint=True, zonal=True)
        self.AddSetting(name="Threshold", comment="Parameters threshold", default=0.5)
        # InObj offset (zonal index of <G>InObj minus global index of G must be constant)
        zs_idx = {s.name: i for i, s in enumerate(self.zonesettings)}
        offs = {zs_idx[g.name + "InObj"] - i for i, g in enumerate(self.globals) if g.name + "InObj" in zs_idx}
        if len(offs) > 1:
            raise ModelError("Not all InObj offsets are the same")
        self.in_obj_offset = offs.pop() if offs else 0
        # derived-setting targets must exist
        names = {s.name for s in self.settings}
        for s in self.settings + self.zonesettings:
            if s.derived is not None and s.derived not in names:
                # the reference generator does not check this (models/multiphase/experimental/d2q9_lee relies on it)
                s.derived, s.equation = None, None
        self.finalized = True
        return self

    def _autosym(self):
        """Option autosym (reference src/conf.R:373-457): every density gets its mirror partner per axis (same group,
        mirrored velocity; `sym=` of AddDensity/AddField overrides, "-name" mirrors with a sign), the reach of every field
        is closed under the mirror maps (x -> -x for autosym=1, a plane through the nodes; x -> 1-x / -1-x for autosym=2,
        a plane on the cell edge), and the node types Symmetry{X,Y,Z}_{plus,minus} (SymmetryEdge... for 2) are added."""
        mode = self.options.get("autosym", 0)
        mode = int(mode) if not isinstance(mode, bool) else (1 if mode else 0)
        self.autosym = mode
        if mode <= 0:
            return
        if mode not in (1, 2):
            raise ModelError("unknown autosym value")
        mirrors = [(-1, 1, 1), (1, -1, 1), (1, 1, -1)]
        by_field = {f.name: f for f in self.fields}
        for g in dict.fromkeys(d.group for d in self.densities):
            D = [d for d in self.densities if d.group == g]
            for d in D:
                v = (d.dx, d.dy, d.dz)
                for ax in range(3):
                    if d.sym[ax] != "":
                        continue
                    if v == (0, 0, 0):
                        partner = d
                    else:
                        sv = tuple(a * b for a, b in zip(v, mirrors[ax]))
                        found = [o for o in D if (o.dx, o.dy, o.dz) == sv]
                        if len(found) == 0:
                            raise ModelError(f"Could not find symmetry for density {d.name}")
                        if len(found) > 1:
                            raise ModelError(f"Too many symmetries for density {d.name}")
                        partner = found[0]
                    d.sym[ax] = partner.name
                    if by_field[d.field].sym[ax] == "":
                        by_field[d.field].sym[ax] = partner.field
        for f in self.fields:
            for ax in range(3):
                if f.sym[ax] == "":
                    f.sym[ax] = f.name
        shift = 0 if mode == 1 else 1
        self.autosym_shift = shift
        self.autosym_name = "Symmetry" if mode == 1 else "SymmetryEdge"
        dirs = {f.name: {(x, y, z) for x in range(f.minx, f.maxx + 1) for y in range(f.miny, f.maxy + 1) for z in range(f.minz, f.maxz + 1)}
                for f in self.fields}
        while True:
            before = {k: set(v) for k, v in dirs.items()}
            for f in self.fields:
                d = dirs[f.name]
                for ax in range(3):
                    nfn = f.sym[ax].lstrip("-")
                    add = set()
                    for p in d:
                        if p[ax] < 0:
                            q = list(p); q[ax] = -p[ax] - shift; add.add(tuple(q))
                        elif p[ax] > 0:
                            q = list(p); q[ax] = -p[ax] + shift; add.add(tuple(q))
                    dirs[nfn] = dirs[nfn] | add
            if dirs == before:
                break
        for f in self.fields:
            pts = dirs[f.name]
            f.minx, f.maxx = min(p[0] for p in pts), max(p[0] for p in pts)
            f.miny, f.maxy = min(p[1] for p in pts), max(p[1] for p in pts)
            f.minz, f.maxz = min(p[2] for p in pts), max(p[2] for p in pts)
        nm = self.autosym_name
        self.AddNodeType(nm + "X_plus", "SYMX"); self.AddNodeType(nm + "X_minus", "SYMX")
        self.AddNodeType(nm + "Y_plus", "SYMY"); self.AddNodeType(nm + "Y_minus", "SYMY")
        if not all(f.minz == 0 and f.maxz == 0 for f in self.fields):
            self.AddNodeType(nm + "Z_plus", "SYMZ"); self.AddNodeType(nm + "Z_minus", "SYMZ")

    def _check_actions(self):
        snames = [s.name for s in self.stages]
        if len(set(snames)) != len(snames):
            raise ModelError("Duplicated Stages' names")
        for a in self.actions:
            if not a.stages:
                raise ModelError(f"There is an empty Action: {a.name}")
            for sn in a.stages:
                if sn not in snames:
                    raise ModelError(f"Some stages in action {a.name} were not defined")
            bufin = [a.name != "Init"] * len(self.fields)
            bufout = [False] * len(self.fields)
            for sn in a.stages:
                st = self.stage(sn)
                ss = [f.save[sn] for f in self.fields]
                sr = [f.read[sn] for f in self.fields]
                loaded = {d.field for d in self.densities if d.load[sn]}
                sl = [f.name in loaded for f in self.fields]
                sr = [False if (not bi and r is None) else r for bi, r in zip(bufin, sr)]
                bad = [f.name for f, bi, r, l in zip(self.fields, bufin, sr, sl) if (not bi) and (bool(r) or l)]
                if bad and not self.permissive_access:
                    raise ModelError(f"Reading fields {bad} in stage '{sn}' that were not yet written in action '{a.name}'")
                bad = [f.name for f, bo, s in zip(self.fields, bufout, ss) if bo and s]
                if bad and not st.can_overwrite and not self.permissive_access:
                    raise ModelError(f"Overwriting fields {bad} in stage '{sn}' written earlier in action '{a.name}'")
                bufout = [bo or s for bo, s in zip(bufout, ss)]
                bufin = list(bufout)
                for f, r in zip(self.fields, sr):
                    f.read[sn] = r
            bad = [f.name for f, bo in zip(self.fields, bufout) if not bo and not f.non_mandatory]
            if bad and not self.permissive_access:
                raise ModelError(f"Fields {bad} were not written in action '{a.name}'")

    def _pack_nodetypes(self):
        """Groups in alphabetical order, ceil(log2(n+1)) bits each, remaining bits = zone."""
        groups: Dict[str, List[NodeType]] = {}
        for n in self.nodetypes:
            groups.setdefault(n.group, []).append(n)
        shift = 0
        packed: List[NodeType] = []
        self.nodetypegroups = []
        for g in sorted(groups):
            tab = groups[g]
            n = len(tab)
            bits = math.ceil(math.log2(n + 1))
            for i, t in enumerate(tab, start=1):
                t.index = i
                t.value = i << shift
                t.mask = ((1 << bits) - 1) << shift
                t.shift = shift
                t.bits = bits
                t.capacity = (1 << bits) - 1
                packed.append(t)
            self.nodetypegroups.append(NodeTypeGroup(g, ((1 << bits) - 1) << shift, shift, bits, (1 << bits) - 1, True))
            shift += bits
        self.flag_bits = 16
        if shift > 14:
            self.flag_bits = 32
            if shift > 30:
                raise ModelError("NodeTypes exceed 32 bits")
        self.zone_bits = self.flag_bits - shift
        self.zone_shift = shift
        self.zone_max = (1 << self.zone_bits) - 1
        packed.append(NodeType("DefaultZone", "SETTINGZONE", 1, 0, self.zone_max << shift, shift,
                               self.zone_bits, self.zone_max, True))
        self.nodetypegroups.append(NodeTypeGroup("SETTINGZONE", self.zone_max << shift, shift, self.zone_bits,
                                                 self.zone_max, True))
        allmax = (1 << self.flag_bits) - 1
        packed.append(NodeType("None", "NONE", 1, 0, 0, 0, 0, 0, False))
        self.nodetypegroups.append(NodeTypeGroup("NONE", 0, 0, 0, 0, False))
        packed.append(NodeType("Clear", "ALL", 1, 0, allmax, 0, self.flag_bits, allmax, False))
        self.nodetypegroups.append(NodeTypeGroup("ALL", allmax, 0, self.flag_bits, allmax, False))
        self.nodetypes = packed

    # ---- convenience ---------------------------------------------------------------------
    def nodetype(self, name) -> NodeType:
        for n in self.nodetypes:
            if n.name == name:
                return n
        raise KeyError(name)

    def bytes_per_update(self, real_size: int) -> int:
        """Reference accounting: 2*nFields*sizeof(real_t)+sizeof(flag_t) (src/main.cpp:127)."""
        return 2 * len(self.fields) * real_size + self.flag_bits // 8

    def summary(self) -> dict:
        return dict(name=self.name, fields=len(self.fields), densities=len(self.densities),
                    settings=[s.name for s in self.settings], zonesettings=[s.name for s in self.zonesettings],
                    globals=[g.name for g in self.globals], quantities=[q.name for q in self.quantities],
                    nodetypes={n.name: n.value for n in self.nodetypes}, stages=[s.name for s in self.stages],
                    actions={a.name: a.stages for a in self.actions}, flag_bits=self.flag_bits,
                    zone_shift=self.zone_shift, zone_max=self.zone_max)
