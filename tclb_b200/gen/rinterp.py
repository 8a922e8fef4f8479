"""Evaluator for the R subset parsed by :mod:`tclb_b200.gen.rlang` + the bridge that lets an unmodified TCLB
``Dynamics.R`` drive :class:`tclb_b200.gen.modeldef.ModelDef` (the reference evaluates it inside ``src/conf.R:362``).
"""
from __future__ import annotations

import math
import os
import re
from collections import OrderedDict
from typing import Any, Callable, Dict, List, Optional

import numpy as np

from .modeldef import ModelDef
from .rlang import NA, NULL, DataFrame, RError, RList, is_str_vec, is_vec, parse, scalar, to_str, truthy, vec


class Env:
    def __init__(self, parent: Optional["Env"] = None):
        self.vars: Dict[str, Any] = {}
        self.parent = parent

    def get(self, name):
        e = self
        while e is not None:
            if name in e.vars:
                return e.vars[name]
            e = e.parent
        raise RError(f"object '{name}' not found")

    def has(self, name):
        e = self
        while e is not None:
            if name in e.vars:
                return True
            e = e.parent
        return False

    def set(self, name, val):
        self.vars[name] = val

    def set_super(self, name, val):
        e = self.parent
        while e is not None:
            if name in e.vars:
                e.vars[name] = val
                return
            if e.parent is None:
                e.vars[name] = val
                return
            e = e.parent
        self.vars[name] = val


class Closure:
    def __init__(self, params, body, env):
        self.params, self.body, self.env = params, body, env


class _Break(Exception):
    pass


class _Next(Exception):
    pass


class _Return(Exception):
    def __init__(self, v):
        self.v = v


class Missing:
    pass


MISSING = Missing()


def _recycle(a, b):
    a, b = vec(a), vec(b)
    if a.ndim == 2 or b.ndim == 2:
        # R recycles the shorter operand down the columns of the matrix (column-major order)
        if a.ndim == 2 and b.ndim == 1 and b.size not in (1, a.size):
            b = np.resize(b, a.size).reshape(a.shape, order="F")
        elif b.ndim == 2 and a.ndim == 1 and a.size not in (1, b.size):
            a = np.resize(a, b.size).reshape(b.shape, order="F")
        elif a.ndim == 2 and b.ndim == 1 and b.size == a.size:
            b = b.reshape(a.shape, order="F")
        elif b.ndim == 2 and a.ndim == 1 and a.size == b.size:
            a = a.reshape(b.shape, order="F")
        return a, b
    na, nb = a.size, b.size
    if na == nb or na == 0 or nb == 0:
        return a, b
    n = max(na, nb)
    if na < n:
        a = np.resize(a, n)
    if nb < n:
        b = np.resize(b, n)
    return a, b


def _num(a):
    a = vec(a)
    if a.dtype == object:
        out = np.array([float("nan") if v is None else float(v) for v in a.ravel()], dtype=float)
        return out.reshape(a.shape)
    if a.dtype == bool:
        return a.astype(np.int64)
    return a


def _simplify_num(x):
    x = np.asarray(x)
    if x.dtype == float and x.size and np.all(np.isfinite(x)) and np.all(x == np.round(x)) and np.all(np.abs(x) < 2 ** 53):
        return x.astype(np.int64)
    return x


class Interp:
    def __init__(self, include_dirs: Optional[List[str]] = None):
        self.globals = Env()
        self.include_dirs = list(include_dirs or [])
        self.out: List[str] = []          # captured cat()/print output (template engine reads it)
        self.sourced: List[str] = []
        # TCLB_PLAIN_C=1: emit C() blocks as written (no reciprocal / shared-product rewriting), for A/B measurements
        self.plain_c = bool(os.environ.get("TCLB_PLAIN_C"))
        self.source_hooks: Dict[str, Callable] = {}
        from .rbuiltins import install
        install(self)
        self.builtins: Dict[str, Any] = dict(self.globals.vars)   # functions stay reachable when a variable takes their name

    # ---- public ----
    def run(self, src: str, env: Optional[Env] = None):
        return self.eval(parse(src), env or self.globals)

    def call(self, fn, args: list, kwargs: "OrderedDict[str, Any]"):
        if isinstance(fn, Closure):
            return self.call_closure(fn, args, kwargs)
        if callable(fn):
            return fn(self, args, kwargs)
        raise RError("attempt to apply non-function")

    def call_closure(self, fn: Closure, args, kwargs):
        env = Env(fn.env)
        names = [p[0] for p in fn.params]
        bound = {}
        extra_pos, extra_kw = [], OrderedDict()
        kw = OrderedDict(kwargs)
        # exact, then partial matching of named arguments
        for k in list(kw):
            if k in names:
                bound[k] = kw.pop(k)
        for k in list(kw):
            cand = [n for n in names if n.startswith(k) and n not in bound and n != "..."]
            if len(cand) == 1 and "..." not in names[:names.index(cand[0])]:
                bound[cand[0]] = kw.pop(k)
        pos = list(args)
        for n in names:
            if n == "...":
                extra_pos = pos
                pos = []
                continue
            if n in bound:
                continue
            if pos:
                bound[n] = pos.pop(0)
        if "..." in names:
            extra_kw = kw
        elif kw:
            raise RError(f"unused argument(s): {', '.join(kw)}")
        elif pos:
            raise RError("unused positional argument")
        for n, default in fn.params:
            if n == "...":
                env.set("...", (extra_pos, extra_kw))
            elif n in bound:
                env.set(n, bound[n])
            elif default is not None:
                env.set(n, ("promise", default, env))
            else:
                env.set(n, MISSING)
        try:
            return self.eval(fn.body, env)
        except _Return as r:
            return r.v

    def force(self, v):
        if isinstance(v, tuple) and len(v) == 3 and v[0] == "promise":
            return self.eval(v[1], v[2])
        return v

    # ---- evaluation ----
    def eval(self, n, env: Env):
        k = n[0]
        if k == "const":
            return n[1]
        if k == "name":
            v = env.get(n[1])
            if isinstance(v, tuple) and len(v) == 3 and v[0] == "promise":
                v = self.eval(v[1], v[2])
                # memoise in the env that holds it
                e = env
                while e is not None and n[1] not in e.vars:
                    e = e.parent
                if e is not None:
                    e.vars[n[1]] = v
            if v is MISSING:
                raise RError(f"argument \"{n[1]}\" is missing, with no default")
            return v
        if k == "paren":
            return self.eval(n[1], env)
        if k == "block":
            v = NULL
            for s in n[1]:
                v = self.eval(s, env)
            return v
        if k == "assign":
            val = self.eval(n[2], env)
            self.assign(n[1], val, env, n[3])
            return val
        if k == "bin":
            return self.binop(n[1], n[2], n[3], env)
        if k == "un":
            v = self.eval(n[2], env)
            if n[1] == "-":
                return self._neg(v)
            if n[1] == "+":
                return v
            return np.logical_not(vec(v).astype(bool)) if is_vec(vec(v)) else (not v)
        if k == "call":
            return self.eval_call(n, env)
        if k == "index":
            return self.index(self.eval(n[1], env), self.eval_index_args(n[2], env))
        if k == "index2":
            return self.index2(self.eval(n[1], env), self.eval(n[2], env))
        if k == "dollar":
            return self.dollar(self.eval(n[1], env), n[2])
        if k == "function":
            return Closure(n[1], n[2], env)
        if k == "if":
            if truthy(self.eval(n[1], env)):
                return self.eval(n[2], env)
            return self.eval(n[3], env) if n[3] is not None else NULL
        if k == "for":
            seq = self.eval(n[2], env)
            items = self.iter_items(seq)
            for it in items:
                env.set(n[1], it)
                try:
                    self.eval(n[3], env)
                except _Break:
                    break
                except _Next:
                    continue
            return NULL
        if k == "while":
            while truthy(self.eval(n[1], env)):
                try:
                    self.eval(n[2], env)
                except _Break:
                    break
                except _Next:
                    continue
            return NULL
        if k == "break":
            raise _Break()
        if k == "next":
            raise _Next()
        if k == "missing":
            return MISSING
        raise RError(f"cannot evaluate node {k}")

    def iter_items(self, seq):
        if seq is NULL:
            return []
        if isinstance(seq, RList):
            return list(seq.values())
        if isinstance(seq, DataFrame):
            return [vec(v) for v in seq.cols.values()]
        seq = vec(seq)
        return [vec(v.item() if hasattr(v, "item") else v) for v in seq.ravel()]

    def _neg(self, v):
        from .poly import Poly, PVec
        if isinstance(v, (Poly, PVec)):
            return -v
        if isinstance(v, DataFrame):
            return DataFrame(OrderedDict((k, [-x for x in c]) for k, c in v.cols.items()))
        return -_num(v)

    def eval_index_args(self, args, env):
        return [(nm, MISSING if a[0] == "missing" else self.eval(a, env)) for nm, a in args]

    def eval_args(self, args, env):
        pos, kw = [], OrderedDict()
        for nm, a in args:
            if a[0] == "missing":
                continue
            if a == ("name", "..."):
                ep, ek = env.get("...")
                pos += [self.force(x) for x in ep]
                for k2, v2 in ek.items():
                    kw[k2] = self.force(v2)
                continue
            v = self.eval(a, env)
            if nm is None:
                pos.append(v)
            else:
                kw[nm] = v
        return pos, kw

    def eval_call(self, n, env):
        fnode, args = n[1], n[2]
        if fnode[0] == "name":
            fname = fnode[1]
            # special forms
            if fname == "missing":
                v = env.get(args[0][1][1])
                return vec(v is MISSING)
            if fname == "quote" or fname == "substitute":
                return ("ast", args[0][1])
            if fname in ("return",):
                raise _Return(self.eval(args[0][1], env) if args else NULL)
            if fname == "invisible":
                return self.eval(args[0][1], env) if args else NULL
            if fname == "exists":
                nm = scalar(self.eval(args[0][1], env))
                return vec(env.has(nm))
            if fname == "function":
                pass
            if fname in ("library", "require", "suppressMessages", "options"):
                return NULL
            if fname == "on.exit":
                return NULL
            if fname == "switch":
                sel = scalar(self.eval(args[0][1], env))
                alts = args[1:]
                if isinstance(sel, str):
                    hit = None
                    for i, (nm, a_) in enumerate(alts):
                        if nm == sel:
                            hit = i
                            break
                    if hit is None:
                        for i, (nm, a_) in enumerate(alts):
                            if nm is None:
                                hit = i
                                break
                    if hit is None:
                        return NULL
                    while hit < len(alts) and alts[hit][1][0] == "missing":      # `a=, b=expr`: fall through
                        hit += 1
                    return self.eval(alts[hit][1], env) if hit < len(alts) else NULL
                i = int(sel) - 1
                return self.eval(alts[i][1], env) if 0 <= i < len(alts) else NULL
            if fname == "stopifnot":
                for _, a in args:
                    if not truthy(np.all(vec(self.eval(a, env)))):
                        raise RError("stopifnot: condition not TRUE")
                return NULL
            if fname == "tryCatch" or fname == "try":
                try:
                    return self.eval(args[0][1], env)
                except RError:
                    return NULL
            if fname == "attach":
                obj = self.eval(args[0][1], env)
                if isinstance(obj, DataFrame):
                    for k2, v2 in obj.cols.items():
                        env.set(k2, vec(list(v2)))
                    self._attached = list(obj.cols)
                return NULL
            if fname == "detach":
                return NULL
            if fname == "assign":
                pos, kw = self.eval_args(args, env)
                self.globals.set(scalar(pos[0]), pos[1])
                return pos[1]
            if fname == "get":
                pos, kw = self.eval_args(args, env)
                return env.get(scalar(pos[0]))
            if fname == "source":
                pos, kw = self.eval_args(args, env)
                return self.source(scalar(pos[0]), env)
            if fname == "sapply" or fname == "lapply" or fname == "Map" or fname == "mapply" or fname == "apply" or fname == "by" or fname == "do.call":
                pass
            fn = env.get(fname) if env.has(fname) else None
            if fn is None:
                raise RError(f"could not find function \"{fname}\"")
            fn = self.force(fn)
            if not callable(fn) and not isinstance(fn, Closure):
                # R looks for a FUNCTION of that name: a variable `C` (a density called C) does not hide C()
                e, fn2 = env, None
                while e is not None and fn2 is None:
                    v = e.vars.get(fname)
                    if v is not None and (callable(v) or isinstance(v, Closure)):
                        fn2 = v
                    e = e.parent
                fn = fn2 or self.builtins.get(fname) or fn
        else:
            fn = self.eval(fnode, env)
        pos, kw = self.eval_args(args, env)
        try:
            return self.call(fn, pos, kw)
        except (_Break, _Next, _Return):
            raise
        except Exception as e:
            if not hasattr(e, "rstack"):
                e.rstack = []
            if len(e.rstack) < 12:
                e.rstack.append(fnode[1] if fnode[0] == "name" else "<fn>")
            raise

    # ---- source() ----
    def find_file(self, name):
        for d in [""] + self.include_dirs:
            p = os.path.join(d, name) if d else name
            if os.path.isfile(p):
                return p
        return None

    def source(self, name, env):
        base = os.path.basename(name)
        key = name[2:] if name.startswith("./") else name
        if key in self.source_hooks:
            return self.source_hooks[key](self, self.globals)
        if base == key and base in self.source_hooks:
            return self.source_hooks[base](self, env)
        p = self.find_file(name)
        if p is None:
            raise RError(f"source: cannot find '{name}' (searched {self.include_dirs})")
        self.sourced.append(p)
        return self.run(open(p).read(), self.globals)

    # ---- operators ----
    def binop(self, op, ln, rn, env):
        from .poly import Poly, PVec
        if op == "&&":
            return vec(truthy(self.eval(ln, env)) and truthy(self.eval(rn, env)))
        if op == "||":
            return vec(truthy(self.eval(ln, env)) or truthy(self.eval(rn, env)))
        a, b = self.eval(ln, env), self.eval(rn, env)
        if isinstance(a, DataFrame) and op != "%in%":      # arithmetic on a data frame works on its matrix of values
            a = np.array([list(c) for c in a.cols.values()], dtype=object).T
            a = _num(a) if all(not isinstance(v, str) for v in a.ravel()) else a
        if isinstance(b, DataFrame) and op != "%in%":
            b = np.array([list(c) for c in b.cols.values()], dtype=object).T
            b = _num(b) if all(not isinstance(v, str) for v in b.ravel()) else b
        if isinstance(a, (Poly, PVec)) or isinstance(b, (Poly, PVec)):
            return self.poly_binop(op, a, b)
        if op == "%in%":
            bs = set(vec(b).ravel().tolist()) if b is not NULL else set()
            return np.array([x in bs for x in vec(a).ravel().tolist()], dtype=bool)
        if op == "%*%":
            A, B = _num(a).astype(float), _num(b).astype(float)
            if A.ndim == 1 and B.ndim == 2:
                A = A.reshape(1, -1) if A.size == B.shape[0] else A.reshape(-1, 1)
            if B.ndim == 1 and A.ndim == 2:
                B = B.reshape(-1, 1) if B.size == A.shape[1] else B.reshape(1, -1)
            if A.ndim == 1 and B.ndim == 1:
                return np.array([[float(A @ B)]])
            return A @ B
        if op == "%o%":
            return np.outer(_num(a), _num(b))
        if op == ":":
            x, y = scalar(a), scalar(b)
            x, y = (int(x) if float(x).is_integer() else x), (int(y) if float(y).is_integer() else y)
            if x <= y:
                return np.arange(x, y + 1)
            return np.arange(x, y - 1, -1)
        if op in ("==", "!="):
            a2, b2 = _recycle(a, b)
            if a2.dtype == object or b2.dtype == object:
                r = np.array([to_str(p) == to_str(q) if (p is not None and q is not None) else False
                              for p, q in zip(a2.ravel().tolist(), b2.ravel().tolist())], dtype=bool)
            else:
                r = a2 == b2
            return r if op == "==" else np.logical_not(r)
        if op in ("<", ">", "<=", ">="):
            a2, b2 = _recycle(a, b)
            if a2.dtype == object or b2.dtype == object:
                pa, pb = [to_str(x) for x in a2.ravel()], [to_str(x) for x in b2.ravel()]
                f = {"<": lambda p, q: p < q, ">": lambda p, q: p > q, "<=": lambda p, q: p <= q, ">=": lambda p, q: p >= q}[op]
                return np.array([f(p, q) for p, q in zip(pa, pb)], dtype=bool)
            a2, b2 = _num(a2), _num(b2)
            return {"<": np.less, ">": np.greater, "<=": np.less_equal, ">=": np.greater_equal}[op](a2, b2)
        if op in ("&", "|"):
            a2, b2 = _recycle(a, b)
            return (np.logical_and if op == "&" else np.logical_or)(a2.astype(bool), b2.astype(bool))
        a2, b2 = _recycle(a, b)
        a2, b2 = _num(a2), _num(b2)
        if op == "+":
            return a2 + b2
        if op == "-":
            return a2 - b2
        if op == "*":
            return a2 * b2
        if op == "/":
            return a2 / np.asarray(b2, dtype=float)
        if op == "^":
            r = np.power(np.asarray(a2, dtype=float), b2)
            return _simplify_num(r) if (a2.dtype != float and b2.dtype != float and np.all(np.asarray(b2) >= 0)) else r
        if op == "%%":
            return np.mod(a2, b2)
        if op == "%/%":
            return np.floor_divide(a2, b2)
        raise RError(f"unsupported operator {op}")

    def poly_binop(self, op, a, b):
        from .poly import Poly, PVec
        def lift(x):
            if isinstance(x, (Poly, PVec)):
                return x
            x = vec(x)
            if x.ndim == 2:
                return x
            if x.size == 1:
                v = scalar(x)
                return Poly.const(float(v)) if not isinstance(v, str) else Poly.sym(v)
            return PVec([float(t) for t in x.tolist()])
        A, B = lift(a), lift(b)
        if op == "%%":          # older polyAlgebra spelling of the symbolic matrix product
            op = "%*%"
        if op == "%*%":
            if isinstance(A, PVec) and isinstance(B, np.ndarray):
                return A @ _num(B).astype(float)
            if isinstance(A, np.ndarray) and isinstance(B, PVec):
                M = _num(A).astype(float)
                return PVec([sum((B[j] * float(M[i, j]) for j in range(M.shape[1])), Poly()) for i in range(M.shape[0])])
            if isinstance(A, PVec) and isinstance(B, PVec):
                return PVec([A @ B])
            raise RError("unsupported %*% operands")
        if op == "^":
            n = int(scalar(b))
            return A ** n if isinstance(A, Poly) else PVec([x ** n for x in A])
        if isinstance(A, Poly) and isinstance(B, PVec):
            A = PVec([A] * len(B))
        if isinstance(B, Poly) and isinstance(A, PVec):
            B = PVec([B] * len(A))
        if isinstance(A, PVec) and isinstance(B, PVec) and len(A) != len(B):
            n = max(len(A), len(B))
            A = PVec([A[i % len(A)] for i in range(n)])
            B = PVec([B[i % len(B)] for i in range(n)])
        if op == "+":
            return A + B
        if op == "-":
            return A - B
        if op == "*":
            return A * B
        if op == "/":
            return A * (B ** -1) if isinstance(B, Poly) else PVec([x * (y ** -1) for x, y in zip(A, B)])
        if op in ("<", ">", "<=", ">="):
            la = [x for x in (A if isinstance(A, PVec) else [A])]
            lb = [x for x in (B if isinstance(B, PVec) else [B])]
            if not all(x.is_const() for x in la + lb):
                raise RError(f"operator {op} needs constant symbolic values")
            f = {"<": lambda p, q: p < q, ">": lambda p, q: p > q, "<=": lambda p, q: p <= q, ">=": lambda p, q: p >= q}[op]
            n = max(len(la), len(lb))
            return np.array([f(float(la[i % len(la)].const_value()), float(lb[i % len(lb)].const_value())) for i in range(n)], dtype=bool)
        if op in ("==", "!="):
            from .poly import to_c
            la = [to_c(x) for x in (A if isinstance(A, PVec) else [A])]
            lb = [to_c(x) for x in (B if isinstance(B, PVec) else [B])]
            r = np.array([p == q for p, q in zip(la, lb)], dtype=bool)
            return r if op == "==" else ~r
        raise RError(f"operator {op} is not defined for symbolic values")

    # ---- indexing ----
    def _pos_index(self, idx, n, names=None):
        """R index -> list of 0-based positions"""
        if idx is MISSING:
            return list(range(n))
        idx = vec(idx)
        if idx.dtype == bool:
            if idx.size < n:
                idx = np.resize(idx, n)
            return [i for i, b in enumerate(idx.tolist()) if b]
        if idx.dtype == object:
            if names is None:
                raise RError("character index on an object without names")
            out = []
            for s in idx.tolist():
                if s not in names:
                    raise RError(f"subscript out of bounds: '{s}'")
                out.append(names.index(s))
            return out
        vals = [int(v) for v in idx.tolist()]
        if vals and all(v <= 0 for v in vals):
            drop = {-v - 1 for v in vals if v != 0}
            return [i for i in range(n) if i not in drop]
        return [v - 1 for v in vals if v != 0]

    def index(self, x, args):
        from .poly import PVec
        idxs = [a for nm, a in args if nm not in ("drop",)]
        drop = True
        for nm, a in args:
            if nm == "drop":
                drop = truthy(a)
        if isinstance(x, PVec):
            pos = self._pos_index(idxs[0], len(x))
            return PVec([x[i] for i in pos])
        if isinstance(x, DataFrame):
            if len(idxs) == 1:
                cols = self._pos_index(idxs[0], x.ncol(), list(x.cols))
                d = DataFrame()
                for c in cols:
                    k = list(x.cols)[c]
                    d.cols[k] = list(x.cols[k])
                return d
            rows = self._pos_index(idxs[0], x.nrow(), getattr(x, "rownames", None))
            cols = self._pos_index(idxs[1], x.ncol(), list(x.cols))
            keys = [list(x.cols)[c] for c in cols]
            if len(keys) == 1 and drop:
                return _col_value([x.cols[keys[0]][r] for r in rows])
            d = DataFrame()
            for k in keys:
                d.cols[k] = [x.cols[k][r] for r in rows]
            if getattr(x, "rownames", None):
                d.rownames = [x.rownames[r] for r in rows]
            return d
        if isinstance(x, RList):
            keys = list(x.keys())
            pos = self._pos_index(idxs[0], len(keys), x.names())
            return RList.make([x[keys[i]] for i in pos], [x.names()[i] for i in pos])
        if x is NULL:
            return NULL
        a = vec(x)
        if a.ndim == 2:
            if len(idxs) == 1:
                flat = a.ravel(order="F")
                pos = self._pos_index(idxs[0], flat.size)
                return flat[pos]
            r = self._pos_index(idxs[0], a.shape[0], getattr(x, "rownames", None))
            c = self._pos_index(idxs[1], a.shape[1], self.colnames(x))
            sub = a[np.ix_(r, c)]
            if drop and (len(r) == 1 or len(c) == 1):
                return sub.ravel()
            return sub
        names = self.names_of(x)
        pos = self._pos_index(idxs[0], a.size, names)
        if len(pos) and max(pos) >= a.size:
            # R: indexing past the end of a vector gives NA, not an error
            ext = np.empty(max(pos) + 1, dtype=object)
            ext[:a.size] = a.tolist()
            ext[a.size:] = None
            a = ext
            names = None
        out = a[pos] if len(pos) else a[:0]
        if names:
            out = NamedVec(out, [names[i] for i in pos])
        return out

    def names_of(self, x):
        return getattr(x, "rnames", None)

    def colnames(self, x):
        return getattr(x, "colnames", None)

    def index2(self, x, i):
        from .poly import PVec
        if isinstance(x, PVec):
            return x[int(scalar(i)) - 1]
        if isinstance(x, RList):
            i = scalar(i)
            if isinstance(i, str):
                if i not in x:
                    return NULL
                return x[i]
            return list(x.values())[int(i) - 1]
        if isinstance(x, DataFrame):
            i = scalar(i)
            return vec(list(x.cols[i] if isinstance(i, str) else list(x.cols.values())[int(i) - 1]))
        a = vec(x)
        i = scalar(i)
        if isinstance(i, str):
            names = self.names_of(x) or []
            return vec(a[names.index(i)])
        return vec(a.ravel(order="F")[int(i) - 1])

    def dollar(self, x, name):
        if isinstance(x, DataFrame):
            if name not in x.cols:
                cand = [k for k in x.cols if k.startswith(name)]
                if len(cand) == 1:
                    name = cand[0]
                else:
                    return NULL
            return _col_value(x.cols[name])
        if isinstance(x, RList):
            if name in x:
                return x[name]
            cand = [k for k in x if k.startswith(name)]
            return x[cand[0]] if len(cand) == 1 else NULL
        if hasattr(x, "r_dollar"):
            return x.r_dollar(name)
        from .poly import Poly
        if isinstance(x, Poly):
            # polyAlgebra keeps a polynomial as a table: one column of exponents per symbol plus the coefficients `.M`
            terms = sorted(x.t.items(), key=lambda kv: (len(kv[0]), kv[0]))
            if name == ".M":
                return np.array([float(c) for _, c in terms], dtype=float)
            return np.array([float(dict(m).get(name, 0)) for m, _ in terms], dtype=float)
        if x is NULL:
            return NULL
        raise RError(f"$ operator is invalid for this value (field {name})")

    # ---- assignment ----
    def assign(self, target, val, env: Env, superassign=False):
        k = target[0]
        if k == "name":
            (env.set_super if superassign else env.set)(target[1], val)
            return
        if k == "const" and is_str_vec(target[1]):
            (env.set_super if superassign else env.set)(scalar(target[1]), val)
            return
        if k in ("index", "index2", "dollar", "call"):
            base = target[1] if k != "call" else target[2][0][1]
            cur = self.eval(base, env) if self._exists(base, env) else NULL
            new = self.replace(target, cur, val, env)
            self.assign(base, new, env, superassign)
            return
        raise RError(f"invalid assignment target ({k}: {str(target)[:120]})")

    def _exists(self, node, env):
        try:
            self.eval(node, env)
            return True
        except RError:
            return False

    def replace(self, target, cur, val, env):
        from .poly import Poly, PVec
        k = target[0]
        if k == "dollar":
            name = target[2]
            if isinstance(cur, DataFrame):
                d = _df_copy(cur)
                from .poly import PVec as _PV
                if isinstance(val, RList):
                    v = list(val.values())
                elif isinstance(val, _PV):
                    v = list(val)
                else:
                    v = vec(val).tolist() if not isinstance(val, list) else val
                n = d.nrow()
                d.cols[name] = [v[i % len(v)] for i in range(n)] if n else list(v)
                return d
            lst = RList(cur) if isinstance(cur, RList) else RList()
            if val is NULL:
                lst.pop(name, None)
            else:
                lst[name] = val
            return lst
        if k == "index2":
            i = scalar(self.eval(target[2], env))
            if isinstance(cur, PVec):
                out = cur.copy()
                out[int(i) - 1] = val if isinstance(val, Poly) else val[0]
                return out
            lst = RList(cur) if isinstance(cur, RList) else RList()
            if isinstance(i, str):
                lst[i] = val
            else:
                keys = list(lst.keys())
                if int(i) - 1 < len(keys):
                    lst[keys[int(i) - 1]] = val
                else:
                    lst[f"\x00{len(keys)}"] = val
            return lst
        if k == "index":
            args = self.eval_index_args(target[2], env)
            idxs = [a for nm, a in args]
            if isinstance(cur, PVec) or isinstance(val, (Poly, PVec)):
                out = cur.copy() if isinstance(cur, PVec) else PVec([float(v) for v in vec(cur).tolist()])
                pos = self._pos_index(idxs[0], len(out))
                vals = list(val) if isinstance(val, PVec) else ([val] if isinstance(val, Poly) else [float(v) for v in vec(val).tolist()])
                for j, p in enumerate(pos):
                    out[p] = vals[j % len(vals)]
                return out
            if isinstance(cur, DataFrame):
                d = _df_copy(cur)
                if len(idxs) == 2:
                    rows = self._pos_index(idxs[0], d.nrow(), getattr(d, "rownames", None))
                    if idxs[1] is not MISSING and is_str_vec(vec(idxs[1])):
                        keys = vec(idxs[1]).tolist()
                        for kk in keys:
                            if kk not in d.cols:
                                d.cols[kk] = [NA] * d.nrow()
                    else:
                        keys = [list(d.cols)[c] for c in self._pos_index(idxs[1], d.ncol(), list(d.cols))]
                    v = [x for col in val.cols.values() for x in col] if isinstance(val, DataFrame) else vec(val).ravel(order="F").tolist()
                    t = 0
                    for kk in keys:
                        for r in rows:
                            d.cols[kk][r] = v[t % len(v)]
                            t += 1
                    return d
                raise RError("unsupported data.frame assignment")
            if isinstance(cur, RList):
                lst = RList(cur)
                keys = list(lst.keys())
                pos = self._pos_index(idxs[0], len(keys), lst.names())
                vals = list(val.values()) if isinstance(val, RList) else [val]
                for j, p in enumerate(pos):
                    lst[keys[p]] = vals[j % len(vals)]
                return lst
            a = np.array(vec(cur), copy=True)
            v = vec(val)
            if a.dtype != object and v.dtype == object:
                a = a.astype(object)
            elif a.dtype in (np.int64, bool) and v.dtype == float:
                a = a.astype(float)
            elif a.dtype == bool and v.dtype == np.int64:
                a = a.astype(np.int64)
            if a.ndim == 2 and len(idxs) == 2:
                r = self._pos_index(idxs[0], a.shape[0])
                c = self._pos_index(idxs[1], a.shape[1], self.colnames(cur))
                vv = np.resize(v, len(r) * len(c)).reshape(len(c), len(r)).T if v.ndim < 2 else v
                a[np.ix_(r, c)] = vv
                return a
            if len(idxs) == 0 or (len(idxs) == 1 and idxs[0] is MISSING):
                # x[] <- v: every element, the shape (and names) of x stay
                flat = a.ravel(order="F")
                flat[:] = np.resize(v.ravel(order="F") if v.ndim > 1 else v, flat.size)
                return flat.reshape(a.shape, order="F") if a.ndim > 1 else flat
            names = self.names_of(cur)
            if idxs[0] is not MISSING and is_str_vec(vec(idxs[0])) and not names:
                names = []
            if names is not None and idxs[0] is not MISSING and is_str_vec(vec(idxs[0])):
                for s in vec(idxs[0]).tolist():
                    if s not in names:
                        names.append(s)
                        a = np.append(a, [None] if a.dtype == object else [0])
                pos = [names.index(s) for s in vec(idxs[0]).tolist()]
            else:
                li = vec(idxs[0]) if idxs[0] is not MISSING else None
                if li is not None and li.dtype == bool and li.size > a.size and a.ndim == 1:
                    a = np.concatenate([a.astype(object), np.array([None] * (li.size - a.size), dtype=object)]) if True else a
                pos = self._pos_index(idxs[0], a.size if a.ndim == 1 else a.size)
            if a.ndim == 2:
                flat = a.ravel(order="F")
                for j, p in enumerate(pos):
                    flat[p] = v.ravel()[j % v.size]
                return flat.reshape(a.shape, order="F")
            mx = max(pos) + 1 if pos else 0
            if mx > a.size:
                a = np.concatenate([a, np.zeros(mx - a.size, dtype=a.dtype) if a.dtype != object else np.array([None] * (mx - a.size), dtype=object)])
            for j, p in enumerate(pos):
                a[p] = v.ravel()[j % v.size]
            return NamedVec(a, names) if names else a
        if k == "call":
            fname = target[1][1]
            if fname == "names":
                nm = vec(val).tolist() if val is not NULL else None
                if isinstance(cur, RList):
                    return RList.make(list(cur.values()), nm)
                if isinstance(cur, DataFrame):
                    d = DataFrame()
                    for kk, v in zip(nm, cur.cols.values()):
                        d.cols[kk] = list(v)
                    return d
                return NamedVec(vec(cur), nm)
            if fname == "class" and isinstance(cur, RList):
                out = RList(cur)
                out.rclass = vec(val).tolist()
                return out
            if fname in ("rownames", "row.names") and isinstance(cur, DataFrame):
                d = _df_copy(cur)
                d.rownames = [to_str(v) for v in vec(val).tolist()] if val is not NULL else None
                return d
            if fname in ("colnames", "rownames", "row.names", "dimnames", "class", "levels"):
                if fname == "colnames" and is_vec(vec(cur)) and vec(cur).ndim == 2:
                    return NamedMat(vec(cur), vec(val).tolist())
                return cur
            if fname == "dim":
                if val is NULL:
                    from .poly import PVec as _PV
                    if isinstance(cur, _PV):
                        cur.dim = None
                        return cur
                    return vec(cur).ravel(order="F") if not isinstance(cur, RList) else cur
                d = [int(t) for t in vec(val).tolist()]
                if isinstance(cur, (RList,)):
                    return cur
                from .poly import PVec as _PV
                if isinstance(cur, _PV):
                    cur.dim = d
                    return cur
                return np.reshape(vec(cur), d, order="F")
            if fname == "attr":
                which = scalar(self.eval(target[2][1][1], env))
                a = AttrArray(vec(cur))
                a.attrs = dict(getattr(cur, "attrs", {}))
                a.attrs[which] = val
                return a
            if fname == "diag":
                a = np.array(vec(cur), dtype=float, copy=True)
                np.fill_diagonal(a, vec(val))
                return a
        raise RError("unsupported replacement form")


def _df_copy(cur):
    d = DataFrame(OrderedDict((kk, list(v)) for kk, v in cur.cols.items()))
    d.rownames = list(cur.rownames) if getattr(cur, "rownames", None) else None
    return d


def _is_cell_scalar(c):
    return c is None or isinstance(c, (str, int, float, bool, np.generic))


def _col_value(col):
    """a data-frame column as an R value: an atomic vector, or a list when the cells are not scalars (I(list(...)))"""
    col = list(col)
    if all(_is_cell_scalar(c) for c in col):
        return vec(col)
    from .poly import Poly, PVec
    if all(isinstance(c, Poly) for c in col):
        return PVec(col)
    return RList.make(col)


class NamedVec(np.ndarray):
    def __new__(cls, arr, names):
        o = np.asarray(arr).view(cls)
        o.rnames = list(names) if names is not None else None
        return o

    def __array_finalize__(self, obj):
        self.rnames = getattr(obj, "rnames", None) if obj is not None and getattr(obj, "shape", None) == self.shape else None


class NamedMat(np.ndarray):
    def __new__(cls, arr, colnames):
        o = np.asarray(arr).view(cls)
        o.colnames = list(colnames)
        return o

    def __array_finalize__(self, obj):
        self.colnames = getattr(obj, "colnames", None)


class AttrArray(np.ndarray):
    def __new__(cls, arr):
        o = np.asarray(arr).view(cls)
        o.attrs = {}
        return o

    def __array_finalize__(self, obj):
        self.attrs = getattr(obj, "attrs", {})
