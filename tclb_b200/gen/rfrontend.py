"""Front-end for unmodified TCLB model directories: evaluates ``Dynamics.R`` with the built-in R-subset interpreter
and fills a :class:`ModelDef` (what ``src/conf.R`` does with real R), then expands ``Dynamics.c.Rt`` with the template
engine (:mod:`tclb_b200.gen.rtemplate`).  ``conf.mk`` is read for OPT/ADJOINT (``src/models.R:28-66``).
"""
from __future__ import annotations

import itertools
import os
import re
from collections import OrderedDict
from typing import Dict, List, Optional

import numpy as np

from .modeldef import ModelDef
from .poly import Poly, PVec
from .rinterp import Env, Interp
from .rlang import NA, NULL, DataFrame, RError, RList, scalar, vec


class _ModelEnv(Env):
    """global environment whose table names (DensityAll, Fields, ...) always reflect the ModelDef being built"""

    def __init__(self, model: ModelDef):
        super().__init__(None)
        self.model = model

    TABLES = ("DensityAll", "Density", "DensityAD", "Fields", "Settings", "ZoneSettings", "Globals", "Quantities", "NodeTypes",
              "Stages", "Actions")

    def get(self, name):
        if name in self.TABLES and name not in self.vars:
            return self.table(name)
        return super().get(name)

    def has(self, name):
        return name in self.TABLES or super().has(name)

    def table(self, name) -> DataFrame:
        m = self.model
        def df(rows: List[dict], keys):
            return DataFrame(OrderedDict((k, [r.get(k) for r in rows]) for k in keys))
        if name in ("DensityAll", "Density", "DensityAD"):
            rows = [dict(name=d.name, field=d.field, dx=d.dx, dy=d.dy, dz=d.dz, comment=d.comment, adjoint=d.adjoint, group=d.group,
                         parameter=d.parameter, average=d.average, default=d.default, nicename=re.sub(r"[\]\[ ]", "", d.name))
                    for d in m.densities if name == "DensityAll" or (d.adjoint == (name == "DensityAD"))]
            return df(rows, ["name", "field", "dx", "dy", "dz", "comment", "adjoint", "group", "parameter", "average", "default", "nicename"])
        if name == "Fields":
            rows = [dict(name=f.name, minx=f.minx, maxx=f.maxx, miny=f.miny, maxy=f.maxy, minz=f.minz, maxz=f.maxz, comment=f.comment,
                         adjoint=f.adjoint, group=f.group, parameter=f.parameter, average=f.average, nicename=re.sub(r"[\]\[ ]", "", f.name))
                    for f in m.fields]
            return df(rows, ["name", "minx", "maxx", "miny", "maxy", "minz", "maxz", "comment", "adjoint", "group", "parameter", "average", "nicename"])
        if name in ("Settings", "ZoneSettings"):
            tab = m.settings if name == "Settings" else m.zonesettings
            rows = [dict(name=s.name, derived=s.derived, equation=s.equation, unit=s.unit, default=s.default, adjoint=s.adjoint,
                         preload=s.preload, comment=s.comment) for s in tab]
            return df(rows, ["name", "derived", "equation", "unit", "default", "adjoint", "preload", "comment"])
        if name == "Globals":
            return df([dict(name=g.name, var=g.var, comment=g.comment, unit=g.unit, adjoint=g.adjoint, op=g.op) for g in m.globals],
                      ["name", "var", "comment", "unit", "adjoint", "op"])
        if name == "Quantities":
            return df([dict(name=q.name, type="vector_t" if q.vector else "real_t", unit=q.unit, adjoint=q.adjoint, vector=q.vector,
                            comment=q.comment) for q in m.quantities], ["name", "type", "unit", "adjoint", "vector", "comment"])
        if name == "NodeTypes":
            return df([dict(name=n.name, group=n.group, value=n.value, mask=n.mask, shift=n.shift, Index="NODE_" + n.name,
                            groupIndex="NODE_" + n.group) for n in m.nodetypes], ["name", "group", "value", "mask", "shift", "Index", "groupIndex"])
        if name == "Stages":
            return df([dict(name=s.name, main=s.main, particle=s.particle) for s in m.stages], ["name", "main", "particle"])
        if name == "Actions":
            return df([dict(name=a.name) for a in m.actions], ["name"])
        raise RError(name)


def _pyval(v):
    """R value -> plain python for ModelDef arguments"""
    if v is NULL:
        return None
    if isinstance(v, RList):
        return [_pyval(x) for x in v.values()]
    a = vec(v)
    vals = [x.item() if hasattr(x, "item") else x for x in a.ravel().tolist()]
    vals = [int(x) if isinstance(x, float) and x.is_integer() and abs(x) < 2 ** 31 else x for x in vals]
    return vals[0] if len(vals) == 1 else vals


def parse_conf_mk(path: str) -> dict:
    out = dict(ADJOINT=0, TEST=True, OPT="")
    if not os.path.isfile(path):
        return out
    for line in open(path):
        line = line.split("#")[0].strip()
        m = re.match(r"^(\w+)\s*=\s*(.*)$", line)
        if m:
            v = m.group(2).strip().strip('"').strip("'")
            out[m.group(1)] = v
    try:
        out["ADJOINT"] = int(out["ADJOINT"])
    except Exception:
        out["ADJOINT"] = 0
    return out


def expand_opt_formula(opt: str) -> List[List[str]]:
    """OPT="a*b+(c)" -> list of option sets, the R-formula semantics the reference uses to enumerate variants
    (terms of the expanded formula: `*` crosses, `+` unions, `:` interaction only)."""
    opt = opt.strip()
    if not opt:
        return [[]]
    toks = re.findall(r"[A-Za-z_][A-Za-z0-9_.]*|[()+*:]", re.sub(r"-\s*1\s*$", "", opt))
    pos = 0

    def parse_sum():
        nonlocal pos
        terms = parse_prod()
        while pos < len(toks) and toks[pos] == "+":
            pos += 1
            for t in parse_prod():
                if t not in terms:
                    terms.append(t)
        return terms

    def parse_prod():
        nonlocal pos
        left = parse_atom()
        while pos < len(toks) and toks[pos] in ("*", ":"):
            op = toks[pos]
            pos += 1
            right = parse_atom()
            inter = [frozenset(a | b) for a in left for b in right]
            if op == "*":
                res = list(left)
                for t in right + inter:
                    if t not in res:
                        res.append(t)
                left = res
            else:
                left = inter
        return left

    def parse_atom():
        nonlocal pos
        t = toks[pos]
        pos += 1
        if t == "(":
            r = parse_sum()
            pos += 1     # ')'
            return r
        return [frozenset([t])]

    minus_one = bool(re.search(r"-\s*1\s*$", opt))
    terms = parse_sum()
    out = [] if minus_one else [[]]
    for t in terms:
        out.append(sorted(t))
    return out


def all_option_names(opt: str) -> List[str]:
    return list(dict.fromkeys(re.findall(r"[A-Za-z_][A-Za-z0-9_.]*", opt)))


def load_dynamics_r(model_dir: str, name: Optional[str] = None, options: Optional[Dict[str, bool]] = None,
                    reference_src: Optional[str] = None, finalize: bool = True, primal_only: bool = True):
    """Evaluate <model_dir>/Dynamics.R and return (ModelDef, Interp).

    `primal_only` (default): a model whose conf.mk says ADJOINT=1 is built as its primal part (the reference's
    `#ifdef ADJOINT` sections left out); `ModelDef.adjoint_requested` keeps what conf.mk asked for."""
    name = name or os.path.basename(os.path.normpath(model_dir))
    conf = parse_conf_mk(os.path.join(model_dir, "conf.mk"))
    opts = {o: False for o in all_option_names(conf.get("OPT", ""))}
    opts["autosym"] = 0
    opts.update(options or {})
    m = ModelDef(name, opts, adjoint=0 if primal_only else conf["ADJOINT"])
    m.adjoint_requested = conf["ADJOINT"]
    I = Interp(include_dirs=[model_dir] + ([reference_src, os.path.join(reference_src, "lib")] if reference_src else []))
    env = _ModelEnv(m)
    env.vars.update(I.globals.vars)
    I.globals = env
    env.set("Options", RList.make([vec(bool(v)) if isinstance(v, bool) else vec(v) for v in opts.values()], list(opts.keys())))
    env.set("MODEL", vec(name))
    env.set("ADJOINT", vec(m.adjoint))
    env.set("include.dir", vec(model_dir))

    import inspect

    def wrap(method, rename=None):
        params = [p for p in inspect.signature(method).parameters.values()]
        pnames = [p.name for p in params if p.kind in (p.POSITIONAL_OR_KEYWORD,)]

        def f(I_, a, k):
            kw = {}
            for kk, v in k.items():
                key = (rename or {}).get(kk, kk)
                if key not in pnames:      # partial matching of argument names, like R
                    cand = [n for n in pnames if n.startswith(key)]
                    if len(cand) == 1:
                        key = cand[0]
                kw[key] = _pyval(v)
            free = [n for n in pnames if n not in kw]
            for x in a:                     # positional arguments fill the parameters not given by name, in order
                if not free:
                    raise RError(f"too many arguments for {method.__name__}")
                kw[free.pop(0)] = _pyval(x)
            method(**kw)
            return NULL
        return f
    env.set("AddDensity", wrap(m.AddDensity))
    env.set("AddField", wrap(m.AddField, {"non.mandatory": "non_mandatory"}))
    env.set("AddSetting", wrap(m.AddSetting))
    env.set("AddGlobal", wrap(m.AddGlobal))
    env.set("AddQuantity", wrap(m.AddQuantity))
    env.set("AddNodeType", wrap(m.AddNodeType))
    env.set("AddDescription", wrap(m.AddDescription))
    env.set("AddAction", wrap(m.AddAction))
    def add_objective(I_, a, k):
        name = _pyval(a[0] if a else k["name"])
        expr = a[1] if len(a) > 1 else k["expr"]
        if isinstance(expr, PVec):
            if len(expr) != 1:
                raise RError("Only one expression for an objective in AddObjective")
            expr = expr[0]
        if not isinstance(expr, Poly):
            raise RError("Objective need to be a polyAlgebra expression")
        m.AddObjective(str(name), expr)
        return NULL
    env.set("AddObjective", add_objective)
    env.set("SetOptions", wrap(m.SetOptions, {"permissive.access": "permissive_access"}))

    def add_stage(I_, a, k):
        kw = {}
        ren = {"load": "load_densities", "save": "save_fields", "read": "read_fields", "load.densities": "load_densities",
               "save.fields": "save_fields", "read.fields": "read_fields", "can.overwrite": "can_overwrite",
               "particle.margin": "particle_margin"}
        for kk, v in k.items():
            key = ren.get(kk, kk)
            pv = _pyval(v)
            if key in ("load_densities", "save_fields", "read_fields") and isinstance(pv, (bool, np.bool_)) is False and not isinstance(pv, list):
                pv = [pv] if isinstance(pv, str) else pv
            kw[key] = pv
        pos = [_pyval(x) for x in a]
        m.AddStage(*pos, **kw)
        return NULL
    env.set("AddStage", add_stage)
    I.source_hooks["conf.R"] = lambda I_, e: NULL
    from . import rlib
    rlib.install(I)
    for scope in (I.globals, env):
        I.builtins.update({k: v for k, v in scope.vars.items() if callable(v)})
    path = os.path.join(model_dir, "Dynamics.R")
    I.run(open(path).read(), env)
    if finalize:
        m.finalize()
    return m, I


def load_model(model_dir: str, name: Optional[str] = None, options: Optional[Dict[str, bool]] = None,
               tclb_root: Optional[str] = None) -> ModelDef:
    """A complete ModelDef (tables + node code) from an unmodified TCLB model directory: ``Dynamics.R`` through the
    R-subset interpreter, ``Dynamics.c.Rt`` through the template engine (or ``Dynamics.c`` as it is).

    `tclb_root` is the root of the source tree the model lives in (needed only for ``<?RT models/... ?>`` includes
    that name files relative to it); by default it is guessed from `model_dir`."""
    from . import poly, rtemplate
    model_dir = os.path.abspath(model_dir)
    if tclb_root is None:
        p = model_dir
        while p != os.path.dirname(p):
            if os.path.isdir(os.path.join(p, "models")) and os.path.isdir(os.path.join(p, "src")):
                tclb_root = p
                break
            p = os.path.dirname(p)
    poly._tmp_counter[0] = 0
    m, I = load_dynamics_r(model_dir, name=name, options=options, reference_src=None, finalize=False)
    search = [model_dir] + ([tclb_root, os.path.join(tclb_root, "src")] if tclb_root else [])
    rt = os.path.join(model_dir, "Dynamics.c.Rt")
    if os.path.exists(rt):
        m.code = rtemplate.render(I, I.globals, rt, search)
    else:
        m.code = rtemplate._read(os.path.join(model_dir, "Dynamics.c"))
    m.finalize()
    if not getattr(m, "tuning", None):
        m.tuning = dict(DIRECTORY_TUNING.get(m.name, {}))
        tj = os.path.join(model_dir, "tuning.json")
        if os.path.exists(tj):
            import json
            m.tuning.update(json.load(open(tj)))
    return m


# Occupancy hints for model directories, by model name (what the native models carry in their own `tuning`): resident
# blocks per SM the one-block-per-tile kernel is compiled for.  Measured on B200, 512^3 fp32, no Globals
# (profiles/r2/ref_lowered_occupancy.txt): d3q27_cumulant 71 registers/7 blocks 25.6 GLUPS -> 64/8 28.2; d3q19 56/9 38.0 ->
# 48/10 40.1.  A directory can override this with a `tuning.json` next to its Dynamics.R.
DIRECTORY_TUNING: Dict[str, Dict[str, int]] = {
    "d3q27_cumulant": {"min_blocks_float": 8},
    "d3q19": {"min_blocks_float": 10},
}

