"""Small symbolic layer used by the model generator.

Plays the role that the external R packages ``polyAlgebra``/``gvector`` play for the
reference generator (used throughout ``/root/reference/src/lib/feq.R`` and
``boundary.R``: ``PV``, ``C``, ``ToC``, ``subst``, ``is.zero``, ``div.mod`` ...).  It is an
independent implementation: a sparse multivariate Laurent polynomial (integer, possibly
negative, exponents) with float/Fraction coefficients, a vector container with numpy
matrix products, and a C expression printer.

Symbols are arbitrary C l-value strings (``"f[3]"``, ``"u.x"``, ``"rho"``).
"""
from __future__ import annotations

from fractions import Fraction
from typing import Dict, Iterable, List, Sequence, Tuple, Union

import os
import numpy as np

Number = Union[int, float, Fraction]
Mono = Tuple[Tuple[str, int], ...]          # sorted ((symbol, power), ...)

_EPS = 1e-14


def _mono_mul(a: Mono, b: Mono) -> Mono:
    if not a:
        return b
    if not b:
        return a
    d = dict(a)
    for s, p in b:
        q = d.get(s, 0) + p
        if q == 0:
            d.pop(s, None)
        else:
            d[s] = q
    return tuple(sorted(d.items()))


def _is_zero_coef(c) -> bool:
    if isinstance(c, Fraction) or isinstance(c, int):
        return c == 0
    return abs(c) < _EPS


class Poly:
    """Sparse Laurent polynomial  sum_k c_k * prod_j s_j^{p_kj}."""

    __slots__ = ("t",)
    __array_priority__ = 1000  # numpy defers to our __r*__ operators

    def __init__(self, terms: Dict[Mono, Number] | None = None):
        self.t: Dict[Mono, Number] = {}
        if terms:
            for m, c in terms.items():
                if not _is_zero_coef(c):
                    self.t[m] = c

    # -- constructors -------------------------------------------------------------------
    @staticmethod
    def sym(name: str) -> "Poly":
        return Poly({((name, 1),): 1})

    @staticmethod
    def const(c: Number) -> "Poly":
        return Poly({(): c})

    @staticmethod
    def lift(x) -> "Poly":
        if isinstance(x, Poly):
            return x
        if isinstance(x, str):
            return Poly.sym(x)
        if isinstance(x, (np.floating, np.integer)):
            x = x.item()
        return Poly.const(x)

    def deriv(self, var: str) -> "Poly":
        """partial derivative with respect to one symbol (Laurent exponents allowed)"""
        out: Dict[Mono, Number] = {}
        for m, c in self.t.items():
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

    # -- queries ------------------------------------------------------------------------
    def is_zero(self) -> bool:
        return not self.t

    def is_const(self) -> bool:
        return all(m == () for m in self.t)

    def const_value(self) -> Number:
        return self.t.get((), 0)

    def symbols(self) -> List[str]:
        out = []
        for m in self.t:
            for s, _ in m:
                if s not in out:
                    out.append(s)
        return out

    def degree_in(self, syms: Iterable[str]) -> int:
        syms = set(syms)
        best = 0
        for m in self.t:
            best = max(best, sum(abs(p) for s, p in m if s in syms))
        return best

    # -- arithmetic ---------------------------------------------------------------------
    def __add__(self, o):
        o = Poly.lift(o)
        r = dict(self.t)
        for m, c in o.t.items():
            r[m] = r.get(m, 0) + c
        return Poly(r)

    __radd__ = __add__

    def __neg__(self):
        return Poly({m: -c for m, c in self.t.items()})

    def __sub__(self, o):
        return self + (-Poly.lift(o))

    def __rsub__(self, o):
        return Poly.lift(o) - self

    def __mul__(self, o):
        if isinstance(o, PVec):
            return NotImplemented
        o = Poly.lift(o)
        r: Dict[Mono, Number] = {}
        for m1, c1 in self.t.items():
            for m2, c2 in o.t.items():
                m = _mono_mul(m1, m2)
                r[m] = r.get(m, 0) + c1 * c2
        return Poly(r)

    __rmul__ = __mul__

    def __pow__(self, n: int):
        n = int(n)
        if n >= 0:
            r = Poly.const(1)
            for _ in range(n):
                r = r * self
            return r
        if len(self.t) != 1:
            raise ValueError("negative power of a non-monomial")
        (m, c), = self.t.items()
        inv = Fraction(1) / c if isinstance(c, (int, Fraction)) else 1.0 / c
        base = Poly({tuple((s, -p) for s, p in m): inv})
        return base ** (-n)

    def __truediv__(self, o):
        o = Poly.lift(o)
        return self * (o ** -1)

    def __rtruediv__(self, o):
        return Poly.lift(o) * (self ** -1)

    # -- transformations ----------------------------------------------------------------
    def subst(self, mapping: Dict[str, "Poly | Number | str"]) -> "Poly":
        """Replace symbols by polynomials (only non-negative powers of replaced symbols
        unless the replacement is a monomial)."""
        mp = {k: Poly.lift(v) for k, v in mapping.items()}
        out = Poly()
        for m, c in self.t.items():
            term = Poly.const(c)
            for s, p in m:
                if s in mp:
                    term = term * (mp[s] ** p)
                else:
                    term = term * Poly({((s, p),): 1})
            out = out + term
        return out

    def truncate(self, syms: Iterable[str], order: int) -> "Poly":
        """Drop terms whose total |power| in ``syms`` exceeds ``order``."""
        syms = set(syms)
        return Poly({m: c for m, c in self.t.items()
                     if sum(abs(p) for s, p in m if s in syms) <= order})

    def split_linear(self, var: str) -> Tuple["Poly", "Poly"]:
        """self == A*var + B with A, B free of var (raises if not linear)."""
        A: Dict[Mono, Number] = {}
        B: Dict[Mono, Number] = {}
        for m, c in self.t.items():
            d = dict(m)
            p = d.pop(var, 0)
            rest = tuple(sorted(d.items()))
            if p == 0:
                B[rest] = B.get(rest, 0) + c
            elif p == 1:
                A[rest] = A.get(rest, 0) + c
            else:
                raise ValueError(f"expression is not linear in {var}")
        return Poly(A), Poly(B)

    def coef(self, mono_syms: Dict[str, int]) -> Number:
        return self.t.get(tuple(sorted(mono_syms.items())), 0)

    def evalf(self, env: Dict[str, float]) -> float:
        tot = 0.0
        for m, c in self.t.items():
            v = float(c)
            for s, p in m:
                v *= env[s] ** p
            tot += v
        return tot

    # -- printing -----------------------------------------------------------------------
    def to_c(self) -> str:
        return to_c(self)

    def __repr__(self):
        return f"Poly({self.to_c()})"


def _fmt_num(c: Number) -> str:
    if isinstance(c, Fraction):
        if c.denominator == 1:
            return f"{c.numerator}.0"
        c = c.numerator / c.denominator
    if isinstance(c, int):
        return f"{c}.0"
    r = repr(float(c))
    if "e" in r or "E" in r:
        mant, ex = r.lower().split("e")
        if "." not in mant:
            mant += ".0"
        return f"{mant}e{int(ex)}"
    if "." not in r and "inf" not in r and "nan" not in r:
        r += ".0"
    return r


def _mono_c(m: Mono) -> Tuple[str, str]:
    num: List[str] = []
    den: List[str] = []
    for s, p in m:
        tgt = num if p > 0 else den
        tgt.extend([s] * abs(p))
    return "*".join(num), "*".join(den)


def _fmt_int(c: Number) -> str:
    f = float(c)
    return str(int(round(f))) if abs(f - round(f)) < 1e-12 else _fmt_num(c)


def to_c(p: "Poly | Number | str", floating: bool = True, wrap_const=None) -> str:
    """C expression text (double literals; the float build rewrites literals later).  `floating=False` prints whole
    numbers as integers (index arithmetic); `wrap_const(text) -> text` wraps every numeric constant (polyAlgebra's
    C(..., float=FALSE, wrap.const=f))."""
    p = Poly.lift(p)
    fmt = _fmt_num if floating else _fmt_int
    if p.is_zero():
        z = fmt(0)
        return wrap_const(z) if wrap_const else z
    parts: List[str] = []
    for m, c in sorted(p.t.items(), key=lambda kv: (len(kv[0]), kv[0])):
        num, den = _mono_c(m)
        neg = (c < 0)
        ac = -c if neg else c
        one = (ac == 1) if isinstance(ac, (int, Fraction)) else abs(ac - 1.0) < 1e-15
        if num:
            body = num if one else f"{num}*{wrap_const(fmt(ac)) if wrap_const else fmt(ac)}"
        elif wrap_const and not den:
            # a bare constant is wrapped with its sign, so that it stays one object of the wrapper's type
            body = wrap_const(fmt(c))
            neg = False if not parts else neg
            if parts:
                body = wrap_const(fmt(ac))
        else:
            body = fmt(ac)
        if den:
            body = f"{body}/({den})" if "*" in den else f"{body}/{den}"
        if not parts:
            parts.append(("-" if neg else "") + body)
        else:
            parts.append((" - " if neg else " + ") + body)
    return "".join(parts)


class PVec:
    """Vector of polynomials with elementwise arithmetic and matrix products."""

    __array_priority__ = 2000

    def __init__(self, items: Iterable):
        self.v: List[Poly] = [Poly.lift(x) for x in items]

    def __len__(self):
        return len(self.v)

    def __iter__(self):
        return iter(self.v)

    def __getitem__(self, i):
        if isinstance(i, (int, np.integer)):
            return self.v[int(i)]
        if isinstance(i, slice):
            return PVec(self.v[i])
        i = np.asarray(i)
        if i.dtype == bool:
            return PVec([x for x, k in zip(self.v, i) if k])
        return PVec([self.v[int(k)] for k in i])

    def __setitem__(self, i, val):
        if isinstance(i, (int, np.integer)):
            self.v[int(i)] = Poly.lift(val)
            return
        idx = np.asarray(i)
        if idx.dtype == bool:
            idx = np.nonzero(idx)[0]
        vals = list(val) if isinstance(val, (PVec, list, tuple)) else [val] * len(idx)
        for k, x in zip(idx, vals):
            self.v[int(k)] = Poly.lift(x)

    def copy(self):
        return PVec(list(self.v))

    def _bin(self, o, f):
        if isinstance(o, (PVec, list, tuple, np.ndarray)):
            o = list(o)
            if len(o) != len(self.v):
                raise ValueError("length mismatch")
            return PVec([f(a, Poly.lift(b)) for a, b in zip(self.v, o)])
        o = Poly.lift(o)
        return PVec([f(a, o) for a in self.v])

    def __add__(self, o): return self._bin(o, lambda a, b: a + b)
    __radd__ = __add__
    def __sub__(self, o): return self._bin(o, lambda a, b: a - b)
    def __rsub__(self, o): return self._bin(o, lambda a, b: b - a)
    def __mul__(self, o): return self._bin(o, lambda a, b: a * b)
    __rmul__ = __mul__
    def __neg__(self): return PVec([-a for a in self.v])

    def __matmul__(self, M):
        """row-vector @ matrix (n,) @ (n,m) -> (m,);  or dot with another PVec."""
        if isinstance(M, PVec):
            return psum(a * b for a, b in zip(self.v, M.v))
        M = np.asarray(M, dtype=float)
        if M.ndim == 1:
            return psum(a * float(c) for a, c in zip(self.v, M) if c != 0)
        assert M.shape[0] == len(self.v), (M.shape, len(self.v))
        out = []
        for j in range(M.shape[1]):
            acc = Poly()
            for i, a in enumerate(self.v):
                c = M[i, j]
                if abs(c) > _EPS:
                    acc = acc + a * _clean(c)
            out.append(acc)
        return PVec(out)

    def subst(self, mapping):
        return PVec([a.subst(mapping) for a in self.v])

    def sum(self):
        return psum(self.v)

    def __repr__(self):
        return "PVec[" + ", ".join(to_c(a) for a in self.v) + "]"


def _clean(c: float) -> Number:
    """Snap floats that are (numerically) small rationals so that printed constants are
    exact (1/3, 1/36 ...) and exactly-cancelling terms really cancel."""
    f = Fraction(c).limit_denominator(5184)
    if abs(float(f) - c) < 1e-13 * max(1.0, abs(c)):
        return f
    return float(c)


def psum(items: Iterable) -> Poly:
    acc = Poly()
    for x in items:
        acc = acc + Poly.lift(x)
    return acc


def PV(*parts) -> PVec:
    """``PV("f[", range(9), "]")`` -> symbols f[0]..f[8]; mirrors the paste-like call
    convention of the reference's symbolic package.  Numbers become constants."""
    n = 1
    seqs = []
    for p in parts:
        if isinstance(p, (str, int, float, Fraction)):
            seqs.append([p])
        else:
            p = list(p)
            seqs.append(p)
            n = max(n, len(p))
    out = []
    for i in range(n):
        toks = [s[i % len(s)] for s in seqs]
        if len(toks) == 1 and not isinstance(toks[0], str):
            out.append(Poly.const(toks[0]))
        else:
            out.append(Poly.sym("".join(str(t) for t in toks)))
    return PVec(out)


_tmp_counter = [0]


def _terms_c(p: Poly, mono_text) -> str:
    """to_c with the text of every monomial supplied by mono_text(mono) -> (numerator text, denominator text)."""
    if p.is_zero():
        return "0.0"
    parts: List[str] = []
    for m, c in sorted(p.t.items(), key=lambda kv: (len(kv[0]), kv[0])):
        num, den = mono_text(m)
        neg = (c < 0)
        ac = -c if neg else c
        one = (ac == 1) if isinstance(ac, (int, Fraction)) else abs(ac - 1.0) < 1e-15
        if num:
            body = num if one else f"{num}*{_fmt_num(ac)}"
        else:
            body = _fmt_num(ac)
        if den:
            body = f"{body}/({den})" if "*" in den else f"{body}/{den}"
        parts.append((("-" if neg else "") if not parts else (" - " if neg else " + ")) + body)
    return "".join(parts)


def c_assign(lhs, rhs, op: str = "=", indent: str = "\t", decl: str = "", optimize: bool = True) -> str:
    """Emit ``lhs = rhs;`` lines (vectorised), the analogue of the reference's ``C()``.

    With ``optimize`` the block is rewritten for the instruction count of the GPU: a division costs ~10 issue
    slots in fp32 (more in fp64), so every symbol that only appears as a divisor gets ONE reciprocal up front and
    the divisions become multiplications, and a product of symbols used by several terms of the block is formed
    once.  Only symbols that the block itself does not assign are touched, so the meaning is unchanged (up to the
    rounding of x/r versus x*(1/r))."""
    if isinstance(lhs, (Poly, str)):
        lhs = [lhs]
    if isinstance(rhs, (Poly, str, int, float, Fraction)):
        rhs = [rhs]
    lhs = list(lhs)
    rhs = list(rhs)
    assert len(lhs) == len(rhs), (len(lhs), len(rhs))
    names = [l if isinstance(l, str) else to_c(l) for l in lhs]
    if not optimize or any(isinstance(r, str) for r in rhs):
        return "\n".join(f"{indent}{decl}{ls} {op} {r if isinstance(r, str) else to_c(r)};" for ls, r in zip(names, rhs)) + "\n"
    polys = [Poly.lift(r) for r in rhs]
    assigned = set(names)
    # ---- reciprocals of divisor symbols the block does not assign ----
    divisors: List[str] = []
    for p in polys:
        for m in p.t:
            for s, e in m:
                if e < 0 and s not in assigned and s not in divisors:
                    divisors.append(s)
    pre: List[str] = []
    rcp: Dict[str, str] = {}
    for s in divisors:
        _tmp_counter[0] += 1
        rcp[s] = f"rcp{_tmp_counter[0]}_"
        pre.append(f"{indent}real_t {rcp[s]}; {rcp[s]} = 1.0/{s};")

    def factors(m: Mono) -> Tuple[List[str], List[str]]:
        num: List[str] = []
        den: List[str] = []
        for s, e in m:
            if e > 0:
                num.extend([s] * e)
            elif s in rcp:
                num.extend([rcp[s]] * (-e))
            else:
                den.extend([s] * (-e))
        return num, den

    # ---- products shared by several terms ----
    uses: Dict[Mono, int] = {}
    for p in polys:
        for m in p.t:
            uses[m] = uses.get(m, 0) + 1
    shared: Dict[Mono, str] = {}
    for m, n in uses.items():
        num, den = factors(m)
        if n >= 2 and not den and len(num) >= 2 and not any(s in assigned for s, _ in m):
            _tmp_counter[0] += 1
            shared[m] = f"mon{_tmp_counter[0]}_"
            pre.append(f"{indent}real_t {shared[m]}; {shared[m]} = {'*'.join(num)};")

    def mono_text(m: Mono) -> Tuple[str, str]:
        if m in shared:
            return shared[m], ""
        num, den = factors(m)
        return "*".join(num), "*".join(den)

    # ---- sums / differences of two terms used by several rows (linear forms such as M %*% f) ----
    if len(polys) >= 2 and os.environ.get("TCLB_LINEAR_CSE", "1") != "0":
        polys = _pair_cse(polys, assigned, factors, mono_text, pre, indent)

    lines = pre + [f"{indent}{decl}{ls} {op} {_terms_c(p, mono_text)};" for ls, p in zip(names, polys)]
    return "\n".join(lines) + "\n"


def _same_coef(a, b) -> bool:
    if isinstance(a, (int, Fraction)) and isinstance(b, (int, Fraction)):
        return a == b
    a, b = float(a), float(b)
    return abs(a - b) <= 1e-14 * max(abs(a), abs(b))


def _pair_cse(polys: List["Poly"], assigned, factors, mono_text, pre: List[str], indent: str) -> List["Poly"]:
    """Greedy pairwise matching over a block of rows (Paar's algorithm for linear transforms, with signs): while two terms
    a, b enter at least two rows with equal (or opposite) coefficients, t = a + b (or a - b) is formed once and the rows use
    t.  A dense moment transform `R = M %*% f` thereby finds the sums and differences of opposite populations on its own:
    d3q19's 19x19 MRT pair falls from ~620 to ~250 additions.  Terms that contain a symbol the block assigns, or a divisor
    that has no reciprocal, stay where they are."""
    rows: List[Dict[Mono, Number]] = [dict(p.t) for p in polys]

    def eligible(m: Mono) -> bool:
        return m != () and not any(s in assigned for s, _ in m) and not factors(m)[1]

    atoms: List[Mono] = []
    for r in rows:
        for m in r:
            if m not in atoms and eligible(m):
                atoms.append(m)
    if len(atoms) < 3:
        return polys
    temp_text: Dict[Mono, str] = {}
    changed = False
    for _ in range(4 * len(atoms) + 64):
        best = None
        for i in range(len(atoms)):
            a = atoms[i]
            ra = [k for k, r in enumerate(rows) if a in r]
            if len(ra) < 2:
                continue
            for j in range(i + 1, len(atoms)):
                b = atoms[j]
                plus = [k for k in ra if b in rows[k] and _same_coef(rows[k][a], rows[k][b])]
                minus = [k for k in ra if b in rows[k] and _same_coef(rows[k][a], -rows[k][b])]
                for sign, ks in ((1, plus), (-1, minus)):
                    if len(ks) >= 2 and (best is None or len(ks) > len(best[3])):
                        best = (a, b, sign, ks)
        if best is None:
            break
        a, b, sign, ks = best
        _tmp_counter[0] += 1
        name = f"lc{_tmp_counter[0]}_"
        t: Mono = ((name, 1),)

        def text(m: Mono) -> str:
            return temp_text[m] if m in temp_text else mono_text(m)[0]

        pre.append(f"{indent}real_t {name}; {name} = {text(a)} {'+' if sign > 0 else '-'} {text(b)};")
        temp_text[t] = name
        for k in ks:
            rows[k][t] = rows[k][a]
            del rows[k][a], rows[k][b]
        atoms.append(t)
        changed = True
    return [Poly(r) for r in rows] if changed else polys
