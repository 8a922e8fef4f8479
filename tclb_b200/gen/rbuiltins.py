"""Builtin functions of the R-subset interpreter (what Dynamics.R / Dynamics.c.Rt files of the reference call)."""
from __future__ import annotations

import os
import math
import re
from collections import OrderedDict

import numpy as np

from .rlang import NA, NULL, DataFrame, RError, RList, is_str_vec, is_vec, scalar, to_str, truthy, vec


def _flat(args):
    out = []
    for a in args:
        if a is NULL:
            continue
        if isinstance(a, RList):
            out += _flat(list(a.values()))
        else:
            out += vec(a).ravel(order="F").tolist()
    return out


def _numv(a):
    from .rinterp import _num
    return _num(a)


def install(I):
    from .rinterp import MISSING, AttrArray, Closure, NamedMat, NamedVec, _simplify_num
    from .poly import PV, Poly, PVec, to_c
    G = I.globals

    def bi(name):
        def deco(f):
            G.set(name, f)
            return f
        return deco

    @bi("c")
    def _c(I, a, k):
        items = list(a) + list(k.values())
        if any(isinstance(x, (Poly, PVec)) for x in items):
            out = []
            for x in items:
                if isinstance(x, PVec):
                    out += list(x)
                elif isinstance(x, Poly):
                    out.append(x)
                else:
                    out += [Poly.const(float(v)) for v in vec(x).tolist()]
            return PVec(out)
        if any(isinstance(x, RList) for x in items):
            vals, names = [], []
            for x in items:
                if isinstance(x, RList):
                    vals += list(x.values()); names += x.names()
                else:
                    vals.append(x); names.append("")
            return RList.make(vals, names)
        names = []
        has_names = bool(k)
        for x in a:
            n = getattr(x, "rnames", None)
            if n:
                has_names = True
                names += n
            else:
                names += [""] * (0 if x is NULL else vec(x).size)
        for kk, x in k.items():
            names += [kk] * vec(x).size
        v = vec(_flat(items))
        return NamedVec(v, names) if has_names else v

    @bi("paste")
    def _paste(I, a, k, default_sep=" "):
        sep = scalar(k.get("sep", vec(default_sep)))
        collapse = k.get("collapse", NULL)
        parts = []
        for x in a:
            if isinstance(x, PVec):
                parts.append([to_c(p) for p in x])
            elif isinstance(x, Poly):
                parts.append([to_c(x)])
            elif isinstance(x, DataFrame):       # a list of columns: each column is one string (as.character of it)
                parts.append([to_str(c[0]) if len(c) == 1 else "c(" + ", ".join(to_str(v) for v in c) + ")" for c in x.cols.values()])
            elif isinstance(x, RList):
                parts.append([to_str(scalar(v)) if not isinstance(v, (RList, DataFrame)) and vec(v).size == 1 else str(v) for v in x.values()])
            else:
                parts.append([to_str(v) for v in vec(x).ravel().tolist()])
        if any(len(p) == 0 for p in parts) and collapse is NULL:
            return np.array([], dtype=object)          # a zero-length argument gives character(0)
        parts = [p for p in parts if len(p) > 0]
        if not parts:
            return vec([""]) if collapse is NULL else vec("")
        n = max(len(p) for p in parts)
        res = [sep.join(p[i % len(p)] for p in parts) for i in range(n)]
        if collapse is not NULL:
            return vec(scalar(collapse).join(res))
        return vec(res)

    G.set("paste0", lambda I, a, k: _paste(I, a, k, default_sep=""))

    @bi("rep")
    def _rep(I, a, k):
        x = a[0]
        times = k.get("times", a[1] if len(a) > 1 else vec(1))
        each = int(scalar(k.get("each", vec(1))))
        if isinstance(x, (Poly, PVec)):
            xs = list(x) if isinstance(x, PVec) else [x]
            xs = [v for v in xs for _ in range(each)]
            return PVec(xs * int(scalar(times)))
        if isinstance(x, RList):
            return RList.make(list(x.values()) * int(scalar(times)))
        xs = [v for v in vec(x).tolist() for _ in range(each)]
        t = vec(times)
        if "length.out" in k:
            n = int(scalar(k["length.out"]))
            return vec([xs[i % len(xs)] for i in range(n)])
        if t.size > 1:
            return vec([v for v, c in zip(xs, t.tolist()) for _ in range(int(c))])
        return vec(xs * int(scalar(t)))

    @bi("seq")
    def _seq(I, a, k):
        if "length.out" in k or "length" in k:
            n = int(scalar(k.get("length.out", k.get("length"))))
            fr = float(scalar(k.get("from", a[0] if a else vec(1))))
            to = float(scalar(k.get("to", a[1] if len(a) > 1 else vec(fr + n - 1))))
            return np.linspace(fr, to, n)
        fr = scalar(k.get("from", a[0] if a else vec(1)))
        to = scalar(k.get("to", a[1] if len(a) > 1 else vec(1)))
        by = scalar(k.get("by", a[2] if len(a) > 2 else vec(1 if to >= fr else -1)))
        if len(a) == 1 and not k:
            return np.arange(1, int(fr) + 1)
        n = int(math.floor((to - fr) / by + 1e-10)) + 1
        return _simplify_num(np.array([fr + i * by for i in range(max(n, 0))], dtype=float))

    G.set("seq_len", lambda I, a, k: np.arange(1, int(scalar(a[0])) + 1))
    G.set("seq_along", lambda I, a, k: np.arange(1, _length(a[0]) + 1))

    def _length(x):
        if x is NULL:
            return 0
        if isinstance(x, (RList, PVec)):
            return len(x)
        if isinstance(x, DataFrame):
            return x.ncol()
        if isinstance(x, Poly):
            return 1
        return vec(x).size

    G.set("length", lambda I, a, k: vec(_length(a[0])))
    G.set("nrow", lambda I, a, k: vec(a[0].nrow()) if isinstance(a[0], DataFrame) else (vec(vec(a[0]).shape[0]) if vec(a[0]).ndim == 2 else NULL))
    G.set("ncol", lambda I, a, k: vec(a[0].ncol()) if isinstance(a[0], DataFrame) else (vec(vec(a[0]).shape[1]) if vec(a[0]).ndim == 2 else NULL))
    G.set("NROW", lambda I, a, k: vec(a[0].nrow()) if isinstance(a[0], DataFrame) else vec(vec(a[0]).shape[0]))
    G.set("dim", lambda I, a, k: vec(list(vec(a[0]).shape)) if not isinstance(a[0], DataFrame) else vec([a[0].nrow(), a[0].ncol()]))

    def _reduce(fn, init=None):
        def f(I, a, k):
            na_rm = truthy(k.get("na.rm", vec(False)))
            if any(isinstance(x, (Poly, PVec)) for x in a):
                from .poly import psum
                items = []
                for x in a:
                    items += list(x) if isinstance(x, PVec) else [x if isinstance(x, Poly) else Poly.const(float(scalar(x)))]
                return psum(items)
            vals = [v for v in _flat(a)]
            if na_rm:
                vals = [v for v in vals if v is not None and not (isinstance(v, float) and math.isnan(v))]
            if not vals:
                return vec(init if init is not None else 0)
            return _simplify_num(np.array([fn(np.array(vals, dtype=float))]))
        return f

    G.set("sum", _reduce(np.sum, 0))
    G.set("prod", _reduce(np.prod, 1))
    G.set("max", _reduce(np.max, float("-inf")))
    G.set("min", _reduce(np.min, float("inf")))
    G.set("mean", _reduce(np.mean))
    G.set("range", lambda I, a, k: vec([float(np.min(_numv(vec(_flat(a))))), float(np.max(_numv(vec(_flat(a)))))]))
    G.set("pmin", lambda I, a, k: np.minimum(*[_numv(x) for x in a[:2]]))
    G.set("pmax", lambda I, a, k: np.maximum(*[_numv(x) for x in a[:2]]))
    for nm, f in dict(abs=np.abs, sqrt=np.sqrt, exp=np.exp, floor=np.floor, ceiling=np.ceil, sin=np.sin, cos=np.cos, tan=np.tan,
                      atan=np.arctan, log2=np.log2, log10=np.log10, sign=np.sign, cumsum=np.cumsum, cumprod=np.cumprod, rev=lambda x: x[::-1]).items():
        G.set(nm, (lambda f: lambda I, a, k: _simplify_num(f(_numv(a[0]).astype(float))) if f not in (np.sqrt, np.exp, np.sin, np.cos, np.tan, np.arctan, np.log2, np.log10)
               else f(_numv(a[0]).astype(float)))(f))
    G.set("rev", lambda I, a, k: (PVec(list(a[0])[::-1]) if isinstance(a[0], PVec) else vec(a[0])[::-1]))
    G.set("log", lambda I, a, k: np.log(_numv(a[0]).astype(float)) / (math.log(float(scalar(a[1]))) if len(a) > 1 else 1.0))
    G.set("round", lambda I, a, k: _simplify_num(np.round(_numv(a[0]).astype(float), int(scalar(a[1])) if len(a) > 1 else int(scalar(k.get("digits", vec(0)))))))
    G.set("as.integer", lambda I, a, k: np.array([int(float(v)) for v in vec(a[0]).ravel().tolist()], dtype=np.int64))
    G.set("as.numeric", lambda I, a, k: np.array([float(v) for v in vec(a[0]).ravel().tolist()], dtype=float))
    G.set("as.double", G.get("as.numeric"))
    G.set("as.character", lambda I, a, k: vec([to_c(p) for p in a[0]]) if isinstance(a[0], PVec) else vec([to_str(v) for v in vec(a[0]).ravel().tolist()]))
    G.set("as.logical", lambda I, a, k: vec(a[0]).astype(bool))
    G.set("as.vector", lambda I, a, k: a[0] if isinstance(a[0], PVec) else vec(a[0]).ravel(order="F"))
    G.set("as.matrix", lambda I, a, k: _as_matrix(a[0]))
    G.set("is.null", lambda I, a, k: vec(a[0] is NULL))
    def _is_na(I, a, k):
        x = a[0]
        isna = lambda v: v is None or (isinstance(v, float) and math.isnan(v))
        if isinstance(x, DataFrame):
            return np.array([[isna(c[i]) for c in x.cols.values()] for i in range(x.nrow())], dtype=bool).reshape(x.nrow(), x.ncol())
        if isinstance(x, (RList, PVec, Poly)):
            return np.zeros(len(x) if not isinstance(x, Poly) else 1, dtype=bool)
        v = vec(x)
        return np.array([isna(t) for t in v.ravel().tolist()], dtype=bool).reshape(v.shape)
    G.set("is.na", _is_na)
    def _is_chr(x):
        if x is NULL or isinstance(x, (RList, DataFrame, PVec, Poly)) or not is_vec(vec(x)):
            return False
        v = vec(x)
        return v.dtype == object and any(isinstance(t, str) for t in v.ravel().tolist())
    def _is_lgl(x):
        if x is NULL or isinstance(x, (RList, DataFrame, PVec, Poly)) or not is_vec(vec(x)):
            return False
        v = vec(x)
        return v.dtype == bool or (v.dtype == object and v.size > 0 and all(t is None or isinstance(t, (bool, np.bool_)) for t in v.ravel().tolist()))
    G.set("is.character", lambda I, a, k: vec(_is_chr(a[0])))
    G.set("is.numeric", lambda I, a, k: vec(is_vec(vec(a[0])) and vec(a[0]).dtype in (np.int64, float)) if not isinstance(a[0], (RList, DataFrame, PVec, Poly))
          else vec(isinstance(a[0], Poly) and a[0].is_const()))      # a constant polynomial counts as a number (polyAlgebra)
    G.set("is.logical", lambda I, a, k: vec(_is_lgl(a[0])))
    G.set("is.function", lambda I, a, k: vec(isinstance(a[0], Closure) or callable(a[0])))
    G.set("is.list", lambda I, a, k: vec(isinstance(a[0], (RList, DataFrame))))
    G.set("is.data.frame", lambda I, a, k: vec(isinstance(a[0], DataFrame)))
    G.set("is.matrix", lambda I, a, k: vec(is_vec(vec(a[0])) and vec(a[0]).ndim == 2))
    G.set("inherits", lambda I, a, k: vec(scalar(a[1]) in ("gvector", "pAlg") and isinstance(a[0], (Poly, PVec))))
    G.set("class", lambda I, a, k: vec(a[0].rclass) if isinstance(a[0], RList) and getattr(a[0], "rclass", None) else vec("gvector" if isinstance(a[0], (Poly, PVec)) else "list" if isinstance(a[0], RList) else "data.frame" if isinstance(a[0], DataFrame)
                                       else "function" if isinstance(a[0], Closure) else "character" if is_str_vec(vec(a[0])) else "numeric"))
    G.set("identical", lambda I, a, k: vec(_flat([a[0]]) == _flat([a[1]])))
    G.set("any", lambda I, a, k: vec(any(bool(v) for v in _flat(a))))
    G.set("all", lambda I, a, k: vec(all(bool(v) for v in _flat(a))))
    G.set("which", lambda I, a, k: np.array([i + 1 for i, v in enumerate(vec(a[0]).ravel(order="F").tolist()) if v], dtype=np.int64))
    G.set("which.max", lambda I, a, k: vec(int(np.argmax(_numv(a[0]))) + 1))
    G.set("which.min", lambda I, a, k: vec(int(np.argmin(_numv(a[0]))) + 1))
    G.set("ifelse", lambda I, a, k: _ifelse(a[0], a[1], a[2]))
    G.set("nchar", lambda I, a, k: vec([len(to_str(v)) for v in vec(a[0]).tolist()]))
    G.set("toupper", lambda I, a, k: vec([to_str(v).upper() for v in vec(a[0]).tolist()]))
    G.set("tolower", lambda I, a, k: vec([to_str(v).lower() for v in vec(a[0]).tolist()]))
    G.set("unique", lambda I, a, k: _unique(a[0]))
    G.set("duplicated", lambda I, a, k: _duplicated(a[0]))
    G.set("setdiff", lambda I, a, k: vec([v for v in dict.fromkeys(vec(a[0]).tolist()) if v not in set(vec(a[1]).tolist())]))
    G.set("intersect", lambda I, a, k: vec([v for v in dict.fromkeys(vec(a[0]).tolist()) if v in set(vec(a[1]).tolist())]))
    G.set("union", lambda I, a, k: vec(list(dict.fromkeys(vec(a[0]).tolist() + vec(a[1]).tolist()))))
    G.set("match", lambda I, a, k: vec([(vec(a[1]).tolist().index(v) + 1) if v in vec(a[1]).tolist() else None for v in vec(a[0]).tolist()]))
    G.set("names", lambda I, a, k: _names(a[0]))
    G.set("colnames", lambda I, a, k: vec(list(a[0].cols)) if isinstance(a[0], DataFrame) else (vec(getattr(a[0], "colnames", None)) if getattr(a[0], "colnames", None) else NULL))
    G.set("rownames", lambda I, a, k: NULL)
    G.set("row.names", lambda I, a, k: NULL)
    G.set("stop", lambda I, a, k: (_ for _ in ()).throw(RError("R stop(): " + "".join(to_str(v) for v in _flat(a)))))
    G.set("warning", lambda I, a, k: NULL)
    G.set("message", lambda I, a, k: NULL)
    G.set("print", lambda I, a, k: NULL)
    G.set("invisible", lambda I, a, k: a[0] if a else NULL)
    G.set("nargs", lambda I, a, k: vec(0))
    G.set("list", lambda I, a, k: RList.make(list(a) + list(k.values()), [""] * len(a) + list(k.keys())))
    G.set("unlist", lambda I, a, k: vec(_flat([a[0]])))
    G.set("I", lambda I, a, k: a[0])
    G.set("t", lambda I, a, k: _t(a[0]))
    G.set("diag", lambda I, a, k: _diag(a, k))
    G.set("solve", lambda I, a, k: np.linalg.inv(_numv(a[0]).astype(float)) if len(a) == 1 else np.linalg.solve(_numv(a[0]).astype(float), _numv(a[1]).astype(float)))
    G.set("chol", lambda I, a, k: np.linalg.cholesky(_numv(a[0]).astype(float)).T)
    G.set("det", lambda I, a, k: vec(float(np.linalg.det(_numv(a[0]).astype(float)))))
    G.set("crossprod", lambda I, a, k: _numv(a[0]).astype(float).T @ _numv(a[1] if len(a) > 1 else a[0]).astype(float))
    G.set("outer", lambda I, a, k: np.outer(_numv(a[0]), _numv(a[1])))
    G.set("rowSums", lambda I, a, k: _simplify_num(_numv(_as_matrix(a[0])).sum(axis=1)))
    G.set("colSums", lambda I, a, k: _simplify_num(_numv(_as_matrix(a[0])).sum(axis=0)))
    G.set("rowMeans", lambda I, a, k: _numv(_as_matrix(a[0])).mean(axis=1))
    G.set("order", lambda I, a, k: _order(a, k))
    G.set("sort", lambda I, a, k: vec(sorted(vec(a[0]).tolist(), reverse=truthy(k.get("decreasing", vec(False))))))
    G.set("table", lambda I, a, k: NULL)
    G.set("nlevels", lambda I, a, k: vec(len(set(vec(a[0]).tolist()))))
    G.set("factor", lambda I, a, k: a[0])
    G.set("numeric", lambda I, a, k: np.zeros(int(scalar(a[0])) if a else 0))
    G.set("integer", lambda I, a, k: np.zeros(int(scalar(a[0])) if a else 0, dtype=np.int64))
    G.set("character", lambda I, a, k: np.array([""] * (int(scalar(a[0])) if a else 0), dtype=object))
    G.set("logical", lambda I, a, k: np.zeros(int(scalar(a[0])) if a else 0, dtype=bool))
    G.set("vector", lambda I, a, k: RList() if scalar(a[0] if a else k.get("mode", vec("list"))) == "list" else np.zeros(int(scalar(a[1] if len(a) > 1 else k.get("length", vec(0))))))
    G.set("nchar", lambda I, a, k: vec([len(to_str(v)) for v in vec(a[0]).tolist()]))
    G.set("substr", lambda I, a, k: vec([to_str(v)[int(scalar(a[1])) - 1:int(scalar(a[2]))] for v in vec(a[0]).tolist()]))
    G.set("strsplit", lambda I, a, k: RList.make([vec(re.split(re.escape(scalar(a[1])) if truthy(k.get("fixed", vec(False))) else scalar(a[1]), to_str(v))) for v in vec(a[0]).tolist()]))
    G.set("trimws", lambda I, a, k: vec([to_str(v).strip() for v in vec(a[0]).tolist()]))
    def _format(I, a, k):
        items = [to_str(v) for v in vec(a[0]).ravel().tolist()]
        just = scalar(k.get("justify", vec("left")))
        if "justify" in k or is_str_vec(vec(a[0])):
            w = max([len(t) for t in items] + [0])
            items = [t.ljust(w) if just == "left" else t.rjust(w) if just == "right" else t.center(w) for t in items]
        return vec(items) if items else np.array([], dtype=object)
    G.set("format", _format)
    G.set("formatC", G.get("format"))
    G.set("file.exists", lambda I, a, k: vec(I.find_file(scalar(a[0])) is not None))
    G.set("readLines", lambda I, a, k: vec(open(I.find_file(scalar(a[0]))).read().split("\n")))
    G.set("Sys.time", lambda I, a, k: vec(0.0))

    # connections: pipe(cmd) + readLines(con) + close(con) (src/conf.R git_version); the command runs in the first
    # include directory's parent; a failing command reads as no lines
    def _pipe(I, a, k):
        r = RList.make([vec(scalar(a[0]))], ["cmd"])
        r.rclass = ["pipe", "connection"]
        return r
    G.set("pipe", _pipe)
    G.set("close", lambda I, a, k: NULL)
    _plain_readlines = G.get("readLines")

    def _readlines(I, a, k):
        if isinstance(a[0], RList) and "pipe" in (getattr(a[0], "rclass", None) or []):
            import subprocess
            try:
                out = subprocess.run(scalar(a[0]["cmd"]), shell=True, capture_output=True, text=True, timeout=20).stdout
            except Exception:
                out = ""
            lines = [l for l in out.split("\n") if l != ""]
            return vec(lines) if lines else np.array([], dtype=object)
        return _plain_readlines(I, a, k)
    G.set("readLines", _readlines)

    def _r_num(x):
        """a number as R's write.table prints it: up to 15 significant digits, no trailing zeros"""
        if isinstance(x, (bool, np.bool_)):
            return "TRUE" if x else "FALSE"
        if isinstance(x, (int, np.integer)):
            return str(int(x))
        if isinstance(x, (float, np.floating)):
            if x != x:
                return "NA"
            t = "%.15g" % float(x)
            return t
        return '"%s"' % str(x).replace('"', '""')

    def _write_table(I, a, k, sep=",", header_quote=True):
        x = a[0] if a else k.get("x")
        file = scalar(a[1]) if len(a) > 1 else scalar(k.get("file"))
        row_names = bool(scalar(k.get("row.names", vec(True))))
        sep = scalar(k.get("sep", vec(sep)))
        if not isinstance(x, DataFrame):
            x = _data_frame(I, [x], OrderedDict())
        cols = list(x.cols.keys())
        n = x.nrow()
        rn = getattr(x, "rownames", None) or [str(i + 1) for i in range(n)]
        lines = [sep.join((['""'] if row_names else []) + ['"%s"' % c for c in cols])]
        for i in range(n):
            cells = [_r_num(x.cols[c][i]) for c in cols]
            lines.append(sep.join((['"%s"' % rn[i]] if row_names else []) + cells))
        path = str(file)
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        with open(path, "w") as f:
            f.write("\n".join(lines) + "\n")
        return NULL
    G.set("write.csv", _write_table)
    G.set("write.table", lambda I, a, k: _write_table(I, a, k, sep=" "))
    G.set("nlevels", lambda I, a, k: vec(0))
    G.set("rows", lambda I, a, k: _rows(a[0]))
    G.set(".GlobalEnv", vec("<globalenv>"))      # assign(..., envir=.GlobalEnv): assign() always writes the global scope here
    G.set("pi", vec(math.pi))
    G.set("T", vec(True))       # ordinary variables in R: models do assign to F and T
    G.set("F", vec(False))
    G.set("letters", vec(list("abcdefghijklmnopqrstuvwxyz")))
    G.set("LETTERS", vec(list("ABCDEFGHIJKLMNOPQRSTUVWXYZ")))

    def _as_matrix(x):
        if isinstance(x, DataFrame):
            cols = [_numv(vec(list(v))) if not is_str_vec(vec(list(v))) else vec(list(v)) for v in x.cols.values()]
            m = np.array([c.tolist() for c in cols]).T
            return NamedMat(_simplify_num(m) if m.dtype != object else m, list(x.cols))
        v = vec(x)
        return v if v.ndim == 2 else v.reshape(-1, 1)

    def _ifelse(c, y, n):
        c = vec(c)
        y, n = vec(y), vec(n)
        out = [(y.ravel()[i % y.size] if b else n.ravel()[i % n.size]) for i, b in enumerate(c.ravel().tolist())]
        r = vec([v.item() if hasattr(v, "item") else v for v in out])
        return r.reshape(c.shape) if c.ndim == 2 else r

    def _unique(x):
        if isinstance(x, DataFrame):
            seen, idx = set(), []
            for i in range(x.nrow()):
                key = tuple(to_str(v[i]) for v in x.cols.values())
                if key not in seen:
                    seen.add(key); idx.append(i)
            return x.subset_rows(idx)
        if isinstance(x, RList):
            return x
        return vec(list(dict.fromkeys(vec(x).ravel().tolist())))

    def _duplicated(x):
        seen, out = set(), []
        for v in vec(x).tolist():
            out.append(v in seen); seen.add(v)
        return np.array(out, dtype=bool)

    def _names(x):
        if isinstance(x, Poly):
            return vec(list(x.symbols()) + [".M"])
        if isinstance(x, PVec) and len(x) == 1:
            return vec(list(x[0].symbols()) + [".M"])
        if isinstance(x, RList):
            return vec(x.names())
        if isinstance(x, DataFrame):
            return vec(list(x.cols))
        n = getattr(x, "rnames", None)
        return vec(n) if n else NULL

    def _t(x):
        from .poly import PVec as _PV
        if isinstance(x, _PV):
            return x
        if isinstance(x, DataFrame):
            x = _as_matrix(x)
        v = vec(x)
        return v.T if v.ndim == 2 else v.reshape(1, -1)

    def _diag(a, k):
        x = a[0] if a else k.get("x", vec(1))
        if "nrow" in k:
            n = int(scalar(k["nrow"]))
            return np.eye(n) * float(scalar(x))
        v = _numv(x)
        if v.ndim == 2:
            return np.diag(v).copy()
        if v.size == 1 and float(v[0]).is_integer() and v[0] >= 1 and len(a) == 1:
            return np.eye(int(v[0]))
        return np.diag(v.astype(float))

    def _order(a, k):
        keys = [vec(x).tolist() for x in a]
        dec = truthy(k.get("decreasing", vec(False)))
        idx = sorted(range(len(keys[0])), key=lambda i: tuple(kk[i] for kk in keys), reverse=dec)
        return np.array([i + 1 for i in idx], dtype=np.int64)

    def _rows(x):
        if x is NULL:
            return RList()
        if isinstance(x, DataFrame):
            return RList.make([x.row(i) for i in range(x.nrow())])
        return x

    @bi("matrix")
    def _matrix(I, a, k):
        data = a[0] if a else k.get("data", vec(NA))
        nrow = k.get("nrow", a[1] if len(a) > 1 else None)
        ncol = k.get("ncol", a[2] if len(a) > 2 else None)
        byrow = truthy(k.get("byrow", a[3] if len(a) > 3 else vec(False)))
        v = vec(data).ravel(order="F")
        if nrow is None and ncol is None:
            nrow, ncol = v.size, 1
        elif nrow is None:
            ncol = int(scalar(ncol)); nrow = -(-v.size // ncol)
        elif ncol is None:
            nrow = int(scalar(nrow)); ncol = -(-v.size // nrow)
        nrow, ncol = int(scalar(vec(nrow))), int(scalar(vec(ncol)))
        v = np.resize(v, nrow * ncol)
        return v.reshape(nrow, ncol) if byrow else v.reshape(nrow, ncol, order="F")

    @bi("array")
    def _array(I, a, k):
        d = [int(t) for t in vec(k.get("dim", a[1] if len(a) > 1 else vec(vec(a[0]).size))).tolist()]
        return np.resize(vec(a[0]).ravel(order="F"), int(np.prod(d))).reshape(d, order="F")

    @bi("expand.grid")
    def _expand_grid(I, a, k):
        cols = [vec(x).tolist() for x in a] + [vec(x).tolist() for x in k.values()]
        names = [f"Var{i + 1}" for i in range(len(a))] + list(k.keys())
        n = int(np.prod([len(c) for c in cols])) if cols else 0
        out = OrderedDict()
        rep = 1
        for nm, c in zip(names, cols):
            out[nm] = [c[(i // rep) % len(c)] for i in range(n)]
            rep *= len(c)
        return DataFrame(out)

    @bi("data.frame")
    def _data_frame(I, a, k):
        cols = OrderedDict()
        for x in a:
            if isinstance(x, DataFrame):
                for kk, v in x.cols.items():
                    cols[kk] = list(v)
            elif is_vec(vec(x)) and vec(x).ndim == 2:
                m = vec(x)
                cn = getattr(x, "colnames", None) or [f"X{i + 1}" for i in range(m.shape[1])]
                for j, kk in enumerate(cn):
                    cols[kk] = m[:, j].tolist()
        for kk, v in k.items():
            if kk in ("stringsAsFactors", "row.names", "check.names"):
                continue
            if isinstance(v, RList):
                cols[kk] = list(v.values())
            elif isinstance(v, PVec):
                cols[kk] = list(v)
            else:
                cols[kk] = vec(v).ravel().tolist() if v is not NULL else []
        return DataFrame(cols)

    @bi("rbind")
    def _rbind(I, a, k):
        items = [x for x in a if x is not NULL]
        if not items:
            return NULL
        if any(isinstance(x, DataFrame) for x in items):
            d = None
            for x in items:
                if not isinstance(x, DataFrame):
                    continue
                d = x if d is None or not d.cols else d.rbind(x)
            return d
        if any(isinstance(x, (Poly, PVec)) for x in items):
            # symbolic scalars stacked into a column (only the one-column case occurs in the models)
            return _c(I, items, OrderedDict())
        rows = [vec(x) if vec(x).ndim == 2 else vec(x).reshape(1, -1) for x in items]
        return np.vstack(rows)

    @bi("cbind")
    def _cbind(I, a, k):
        items = [(None, x) for x in a if x is not NULL] + [(kk, x) for kk, x in k.items() if x is not NULL]
        if not items:
            return NULL
        if any(isinstance(x, DataFrame) for _, x in items):
            cols = OrderedDict()
            for kk, x in items:
                if isinstance(x, DataFrame):
                    for c, v in x.cols.items():
                        cols[c] = list(v)
                else:
                    cols[kk or f"V{len(cols)+1}"] = vec(x).tolist()
            return DataFrame(cols)
        n = max(vec(x).shape[0] for _, x in items)
        mats, names = [], []
        for kk, x in items:
            v = vec(x)
            if v.ndim == 1:
                v = np.resize(v, n).reshape(-1, 1)
                names.append(kk or "")
            else:
                names += getattr(x, "colnames", None) or [""] * v.shape[1]
            mats.append(v)
        m = np.hstack(mats)
        return NamedMat(m, names) if any(names) else m

    @bi("merge")
    def _merge(I, a, k):
        x, y = a[0], a[1]
        if not isinstance(x, DataFrame):
            x = _data_frame(I, [x], OrderedDict())
        if not isinstance(y, DataFrame):
            y = _data_frame(I, [y], OrderedDict())
        by = vec(k["by"]).tolist() if "by" in k else [c for c in x.cols if c in y.cols]
        out = OrderedDict((c, []) for c in by)
        xo = [c for c in x.cols if c not in by]
        yo = [c for c in y.cols if c not in by]
        def nm(c, suf):
            return c + suf if (c in xo and c in yo) else c
        for c in xo:
            out[nm(c, ".x")] = []
        for c in yo:
            out[nm(c, ".y")] = []
        rows = []
        for i in range(x.nrow()):
            for j in range(y.nrow()):
                if all(to_str(x.cols[c][i]) == to_str(y.cols[c][j]) for c in by):
                    rows.append((i, j))
        rows.sort(key=lambda ij: tuple(x.cols[c][ij[0]] for c in by))
        for i, j in rows:
            for c in by:
                out[c].append(x.cols[c][i])
            for c in xo:
                out[nm(c, ".x")].append(x.cols[c][i])
            for c in yo:
                out[nm(c, ".y")].append(y.cols[c][j])
        return DataFrame(out)

    def _apply_fn(I, fn, *args):
        return I.call(fn, list(args), OrderedDict())

    @bi("sapply")
    def _sapply(I, a, k):
        X, fn = a[0], a[1] if len(a) > 1 else k["FUN"]
        items = I.iter_items(X)
        res = [I.call(fn, [it] + list(a[2:]), OrderedDict((kk, v) for kk, v in k.items() if kk not in ("FUN", "simplify", "USE.NAMES"))) for it in items]
        if all(is_vec(vec(r)) and not isinstance(r, (RList, DataFrame, Poly, PVec)) and vec(r).size == 1 for r in res) and res:
            return vec([scalar(r) for r in res])
        if res and all(is_vec(vec(r)) and not isinstance(r, (RList, DataFrame, Poly, PVec)) for r in res) and len({vec(r).size for r in res}) == 1:
            return np.array([vec(r).tolist() for r in res]).T
        return RList.make(res)

    @bi("lapply")
    def _lapply(I, a, k):
        X, fn = a[0], a[1] if len(a) > 1 else k["FUN"]
        names = X.names() if isinstance(X, RList) else None
        return RList.make([I.call(fn, [it] + list(a[2:]), OrderedDict()) for it in I.iter_items(X)], names)

    @bi("do.call")
    def _do_call(I, a, k):
        fn = a[0]
        if is_str_vec(vec(fn)) if not isinstance(fn, Closure) and not callable(fn) else False:
            fn = I.globals.get(scalar(fn))
        args = a[1]
        pos, kw = [], OrderedDict()
        if isinstance(args, RList):
            for nm, v in zip(args.names(), args.values()):
                if nm:
                    kw[nm] = v
                else:
                    pos.append(v)
        elif isinstance(args, DataFrame):      # a data frame is a list of its columns
            from .rinterp import _col_value
            for nm, col in args.cols.items():
                if isinstance(fn, Closure) and nm in [p_[0] for p_ in fn.params]:
                    kw[nm] = _col_value(col)
                else:
                    pos.append(_col_value(col))
        return I.call(fn, pos, kw)

    @bi("apply")
    def _apply(I, a, k):
        X, margin, fn = _as_matrix(a[0]), int(scalar(a[1])), a[2]
        res = []
        n = X.shape[0] if margin == 1 else X.shape[1]
        for i in range(n):
            v = X[i, :] if margin == 1 else X[:, i]
            res.append(I.call(fn, [np.asarray(v)] + list(a[3:]), OrderedDict()))
        if all(vec(r).size == 1 for r in res):
            return vec([scalar(r) for r in res])
        return np.array([vec(r).tolist() for r in res]).T

    @bi("by")
    def _by(I, a, k):
        data, idx, fn = a[0], vec(a[1]).tolist(), a[2]
        out = []
        for g in sorted(dict.fromkeys(idx), key=lambda v: to_str(v)):
            rows = [i for i, v in enumerate(idx) if v == g]
            out.append(I.call(fn, [data.subset_rows(rows)], OrderedDict()))
        return RList.make(out)

    @bi("Reduce")
    def _Reduce(I, a, k):
        fn, items = a[0], I.iter_items(a[1])
        acc = items[0]
        for it in items[1:]:
            acc = I.call(fn, [acc, it], OrderedDict())
        return acc

    @bi("Filter")
    def _Filter(I, a, k):
        return RList.make([it for it in I.iter_items(a[1]) if truthy(I.call(a[0], [it], OrderedDict()))])

    @bi("mapply")
    def _mapply(I, a, k):
        fn = a[0] if a else k["FUN"]
        lists = [I.iter_items(x) for x in a[1:]]
        n = max(len(l) for l in lists)
        return RList.make([I.call(fn, [l[i % len(l)] for l in lists], OrderedDict()) for i in range(n)])

    @bi("sprintf")
    def _sprintf(I, a, k):
        fmt = to_str(scalar(a[0]))
        args = [vec(x).ravel(order="F").tolist() for x in a[1:]]
        n = max([len(x) for x in args] + [1])
        out = []
        for i in range(n):
            vals = []
            for x in args:
                v = x[i % len(x)]
                vals.append(v)
            f2 = re.sub(r"%(\d*)l([dfge])", r"%\1\2", fmt)
            conv = re.findall(r"%[-+ 0#]*\d*(?:\.\d+)?([dfgeEsxXi])", f2)
            vals2 = []
            for v, c in zip(vals, conv):
                vals2.append(int(v) if c in "dixX" else float(v) if c in "fgeE" else to_str(v))
            out.append(f2 % tuple(vals2))
        return vec(out)

    def _regex(pat, fixed):
        return re.escape(pat) if fixed else pat.replace("[[:space:]]", r"\s").replace("[[:alnum:]]", r"[A-Za-z0-9]").replace("[[:digit:]]", r"\d")

    G.set("grepl", lambda I, a, k: np.array([re.search(_regex(scalar(a[0]), truthy(k.get("fixed", vec(False)))), to_str(v)) is not None for v in vec(a[1]).tolist()], dtype=bool))
    G.set("grep", lambda I, a, k: np.array([i + 1 for i, v in enumerate(vec(a[1]).tolist()) if re.search(_regex(scalar(a[0]), False), to_str(v))], dtype=np.int64))
    G.set("sub", lambda I, a, k: vec([re.sub(_regex(scalar(a[0]), truthy(k.get("fixed", vec(False)))), scalar(a[1]).replace("\\\\", "\\"), to_str(v), count=1) for v in vec(a[2]).tolist()]))
    G.set("gsub", lambda I, a, k: vec([re.sub(_regex(scalar(a[0]), truthy(k.get("fixed", vec(False)))), scalar(a[1]).replace("\\\\", "\\"), to_str(v)) for v in vec(a[2]).tolist()]))

    @bi("cat")
    def _cat(I, a, k):
        sep = scalar(k.get("sep", vec(" ")))
        parts = []
        for x in a:
            if isinstance(x, PVec):
                parts += [to_c(p) for p in x]
            elif isinstance(x, Poly):
                parts.append(to_c(x))
            else:
                parts += [to_str(v) for v in vec(x).ravel().tolist()]
        I.out.append(sep.join(parts))
        return NULL

    G.set("writeLines", lambda I, a, k: (I.out.append("\n".join(to_str(v) for v in vec(a[0]).tolist()) + "\n"), NULL)[1])
    G.set("attr", lambda I, a, k: getattr(a[0], "attrs", {}).get(scalar(a[1]), NULL))
    G.set("nlevels", lambda I, a, k: vec(0))
    G.set("head", lambda I, a, k: vec(a[0])[:int(scalar(a[1] if len(a) > 1 else vec(6)))])
    G.set("tail", lambda I, a, k: vec(a[0])[-int(scalar(a[1] if len(a) > 1 else vec(6))):])
    G.set("max.col", lambda I, a, k: NULL)
    G.set("numeric_version", lambda I, a, k: a[0])
    G.set("Recall", lambda I, a, k: NULL)
    install_poly(I)


def install_poly(I):
    """polyAlgebra / gvector surface used by model files: PV, C, ToC, subst, is.zero, V, gapply-free helpers."""
    from collections import OrderedDict
    from .poly import PV, Poly, PVec, c_assign, to_c
    from .rinterp import Closure
    G = I.globals

    def _pv(I, a, k):
        parts = []
        for x in a:
            if isinstance(x, PVec):
                parts.append([to_c(p) for p in x])
            elif isinstance(x, Poly):
                parts.append([to_c(x)])
            else:
                v = vec(x)
                parts.append([t.item() if hasattr(t, "item") else t for t in v.ravel(order="F").tolist()])
        n = max(len(p) for p in parts) if parts else 0
        out = []
        for i in range(n):
            toks = [p[i % len(p)] for p in parts]
            if len(toks) == 1 and toks[0] is None:
                out.append(Poly.sym("NA"))
            elif len(toks) == 1 and not isinstance(toks[0], str):
                out.append(Poly.const(float(toks[0])))
            else:
                out.append(Poly.sym("".join(to_str(t) for t in toks)))
        return PVec(out)
    G.set("PV", _pv)
    def _V(I, a, k):
        """gvector V(...): every argument, flattened, as one symbolic vector"""
        out = []
        for x in a:
            out += list(x) if isinstance(x, PVec) else [x] if isinstance(x, Poly) else [Poly.const(float(v)) for v in vec(x).ravel(order="F").tolist()]
        return PVec(out)
    G.set("V", _V)

    def _toc(I, a, k):
        x = a[0]
        if isinstance(x, PVec):
            return vec([to_c(p) for p in x])
        if isinstance(x, Poly):
            return vec(to_c(x))
        return vec([to_str(v) for v in vec(x).tolist()])
    G.set("ToC", _toc)

    def _C(I, a, k):
        eq = scalar(k.get("eq", vec("=")))
        sep = scalar(k.get("sep", vec(";\n")))
        floating = truthy(k["float"]) if "float" in k else True
        wrap = None
        if "wrap.const" in k and k["wrap.const"] is not NULL:
            wfn = k["wrap.const"]
            wrap = lambda text: to_str(scalar(I.call(wfn, [vec(float(text))], OrderedDict())))
        if not floating or wrap:
            _base = to_c
            def to_c_(p_):
                return _base(p_, floating=floating, wrap_const=wrap)
        else:
            to_c_ = to_c
        if len(a) == 1:
            x = a[0]
            items = [to_c_(p) for p in x] if isinstance(x, PVec) else [to_c_(x) if isinstance(x, Poly) else to_str(scalar(x))]
            I.out.append((sep if "sep" in k else ", ").join(items) if len(items) > 1 else items[0])
            return NULL
        lhs, rhs = a[0], a[1]
        if (isinstance(lhs, PVec) and isinstance(rhs, PVec) and len(lhs) == len(rhs) >= 2 and not getattr(I, "plain_c", False)
                and eq in ("=", "+=", "-=") and all(len(p.t) == 1 for p in lhs)):
            # symbolic block: same rewriting as the native models get (one reciprocal per divisor, shared products)
            I.out.append(c_assign([to_c_(p) for p in lhs], list(rhs), op=eq, indent=""))
            return NULL
        L = [to_c_(p) for p in lhs] if isinstance(lhs, PVec) else [to_c_(lhs)] if isinstance(lhs, Poly) else [to_str(v) for v in vec(lhs).tolist()]
        if isinstance(rhs, PVec):
            R = [to_c_(p) for p in rhs]
        elif isinstance(rhs, Poly):
            R = [to_c_(rhs)]
        else:
            R = [to_c_(Poly.const(float(v))) if not isinstance(v, str) else v for v in vec(rhs).ravel(order="F").tolist()]
        n = max(len(L), len(R))
        I.out.append("".join(f"{L[i % len(L)]} {eq} {R[i % len(R)]};\n" for i in range(n)))
        return NULL
    G.set("C", _C)

    def _subst(I, a, k):
        x = a[0]
        mp = {}
        for kk, v in k.items():
            mp[kk] = v if isinstance(v, Poly) else (v[0] if isinstance(v, PVec) else Poly.const(float(scalar(v))) if not isinstance(scalar(v), str) else Poly.sym(scalar(v)))
        for extra in a[1:]:
            if isinstance(extra, RList):
                for kk, v in zip(extra.names(), extra.values()):
                    mp[kk] = v if isinstance(v, Poly) else v[0] if isinstance(v, PVec) else Poly.const(float(scalar(v)))
        return x.subst(mp) if isinstance(x, (Poly, PVec)) else x
    G.set("subst", _subst)
    G.set("is.zero", lambda I, a, k: (np.array([p.is_zero() for p in a[0]], dtype=bool) if isinstance(a[0], PVec) else vec(a[0].is_zero()) if isinstance(a[0], Poly)
                                      else vec(a[0]) == 0))

    def _gapply(I, a, k):
        """gvector::gapply: apply FUN to every element of a symbolic vector; simplify=TRUE returns a plain vector"""
        x, fn = a[0], a[1]
        simp = truthy(k.get("simplify", vec(False)))
        items = list(x) if isinstance(x, PVec) else [x]
        res = [I.call(fn, [it], OrderedDict()) for it in items]
        if simp:
            flat = []
            for r in res:
                r = vec(r)
                flat.extend(r.ravel().tolist() if r.size else [0.0])
            return vec(flat)
        if all(isinstance(r, Poly) for r in res):
            return PVec(res)
        return RList.make(res)
    G.set("gapply", _gapply)

    def _append(I, a, k):
        x, values = a[0], a[1]
        after = int(scalar(k["after"])) if "after" in k else None
        if isinstance(x, RList) or isinstance(values, RList):
            xs = list(x.items()) if isinstance(x, RList) else [(f"\x00a{i}", v) for i, v in enumerate(_flat([x]))]
            vs = list(values.items()) if isinstance(values, RList) else [(f"\x00b{i}", v) for i, v in enumerate(_flat([values]))]
            allv = xs + vs if after is None else xs[:after] + vs + xs[after:]
            return RList.make([v for _, v in allv], [n if not n.startswith("\x00") else "" for n, _ in allv])
        xs, vs = _flat([x]) if x is not NULL else [], _flat([values]) if values is not NULL else []
        return vec(xs + vs if after is None else xs[:after] + vs + xs[after:])
    G.set("append", _append)

    def _replace(I, a, k):
        x = vec(a[0]).copy()
        idx = vec(a[1])
        vals = vec(a[2])
        flat = x.ravel(order="F")
        pos = np.nonzero(idx.ravel(order="F"))[0] if idx.dtype == bool else (idx.astype(int) - 1)
        for j, p_ in enumerate(pos.tolist()):
            flat[p_] = vals.ravel()[j % vals.size]
        out = flat.reshape(x.shape, order="F")
        cn = getattr(a[0], "colnames", None)
        if cn and out.ndim == 2:
            from .rinterp import NamedMat
            out = NamedMat(out, cn)
        return out
    G.set("replace", _replace)
    def _provide_dimnames(I, a, k):
        from .rinterp import NamedMat
        base = k.get("base", a[1] if len(a) > 1 else None)
        if isinstance(base, RList) and len(base) >= 2 and vec(a[0]).ndim == 2:
            cn = [to_str(v) for v in vec(list(base.values())[1]).tolist()]
            m = vec(a[0])
            return NamedMat(m, (cn * (m.shape[1] // max(1, len(cn)) + 1))[:m.shape[1]])
        return a[0]
    G.set("provideDimnames", _provide_dimnames)
    G.set("as.data.frame", lambda I, a, k: I.call(G.get("data.frame"), [a[0]], OrderedDict()))
    def _deriv_poly(p, var):
        out = {}
        for m, c in p.t.items():
            d = dict(m)
            e = d.get(var, 0)
            if e == 0:
                continue
            if e == 1:
                d.pop(var)
            else:
                d[var] = e - 1
            key = tuple(sorted(d.items()))
            out[key] = out.get(key, 0) + c * e
        return Poly(out)

    def _deriv(I, a, k):
        """polyAlgebra deriv(expr, name): partial derivative of a (Laurent) polynomial with respect to one symbol"""
        x, var = a[0], scalar(a[1])
        if isinstance(x, PVec):
            return PVec([_deriv_poly(p_, var) for p_ in x])
        if isinstance(x, Poly):
            return _deriv_poly(x, var)
        return Poly.const(0.0)
    G.set("deriv", _deriv)
    G.set("c_header", lambda I, a, k: NULL)
    G.set("c_table_decl", lambda I, a, k: a[0])
    G.set("ifdef", lambda I, a, k: NULL)
