"""Generator-time LBM helpers: velocity sets, moment bases, equilibria, boundary closures,
cumulant transforms.  Independent re-implementation of the capabilities the reference keeps
in ``src/lib/feq.R`` (MRT_polyMatrix/MRT_eq, :9-113), ``src/lib/boundary.R`` (Bounce,
FullBounceBack, Symmetry, ZouHe, ZouHeRewrite, :2-198), ``src/lib/lattice.R`` and
``src/lib/cumulant.R``.  Everything returns C text built with :mod:`tclb_b200.gen.poly`.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from itertools import product
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

from .poly import PV, Poly, PVec, c_assign, psum, to_c

# ------------------------------------------------------------------------------------------
# velocity sets
# ------------------------------------------------------------------------------------------

def d2q9_velocities() -> np.ndarray:
    return np.array([[0, 0], [1, 0], [0, 1], [-1, 0], [0, -1],
                     [1, 1], [-1, 1], [-1, -1], [1, -1]], dtype=int)


def tensor_velocities(dim: int) -> np.ndarray:
    """(0,1,-1)^dim with the first coordinate fastest (expand.grid order)."""
    x = [0, 1, -1]
    rows = [tuple(reversed(t)) for t in product(x, repeat=dim)]
    return np.array(rows, dtype=int)


def tensor_names(prefix: str, dim: int) -> List[str]:
    """names like f012 : digit = 0,1,2 for velocity 0,+1,-1, first coordinate first."""
    rows = [tuple(reversed(t)) for t in product(range(3), repeat=dim)]
    return [prefix + "".join(str(d) for d in r) for r in rows]


# ------------------------------------------------------------------------------------------
# polynomial moment basis and equilibrium
# ------------------------------------------------------------------------------------------

@dataclass
class MomentBasis:
    order: np.ndarray          # order of each moment (column)
    mat: np.ndarray            # (Q, Q): column j = moment polynomial j evaluated on velocities
    p: Optional[np.ndarray]    # exponents (Q, d) when the basis is the raw monomial one


def poly_moment_matrix(U: np.ndarray) -> MomentBasis:
    """Raw monomial moments prod_k c_k^{p_k} with exponent vectors taken from the velocity
    set itself (-1 -> 2), ordered by total order (stable)."""
    U = np.asarray(U, dtype=int)
    if np.any(np.abs(U) > 1):
        raise ValueError("only nearest-neighbour velocity sets are supported")
    p = np.where(U < 0, 2, U)
    idx = np.argsort(p.sum(axis=1), kind="stable")
    p = p[idx]
    Q = U.shape[0]
    W = np.zeros((Q, Q))
    for j in range(Q):
        W[:, j] = np.prod(U.astype(float) ** p[j], axis=1)
    return MomentBasis(order=p.sum(axis=1), mat=W, p=p)


@dataclass
class Equilibrium:
    U: np.ndarray
    Req: PVec                  # equilibrium moments in basis `mat`
    mat: np.ndarray
    order: np.ndarray
    feq: PVec
    rho: Poly
    J: PVec


def integer_orthogonalise(M: np.ndarray) -> np.ndarray:
    """Gram-Schmidt on columns keeping integer entries (scaled, not normalised)."""
    from math import gcd
    M = np.array(M, dtype=np.int64)
    for i in range(1, M.shape[1]):
        a = (M[:, :i].T @ M[:, i]).astype(np.int64)
        b = (M[:, :i] ** 2).sum(axis=0).astype(np.int64)
        g = np.array([gcd(int(x), int(y)) for x, y in zip(a, b)], dtype=np.int64)
        g[g == 0] = 1
        a //= g
        b //= g
        l = 1
        for x in b:
            l = l * int(x) // gcd(l, int(x))
        M[:, i] = M[:, i] * l - M[:, :i] @ (l * a // b)
    return M.astype(float)


def mrt_eq(U, rho="rho", J=("Jx", "Jy", "Jz"), sigma2=1.0 / 3.0, order: int = 2,
           orthogonal: bool = True, mat: Optional[np.ndarray] = None, correction=None) -> Equilibrium:
    """Equilibrium moments: rho * prod_k {1, J_k/rho, J_k^2/rho^2 + sigma2}, truncated at
    `order` in J, expressed either in the raw monomial basis, an integer-orthogonalised one
    or a user matrix `mat`."""
    U = np.asarray(U, dtype=int)
    d = U.shape[1]
    rho_p = Poly.lift(rho)
    Jv = PVec(list(J)[:d]) if not isinstance(J, PVec) else PVec(J.v[:d])
    Jnames = []
    for j in Jv:
        Jnames.extend(j.symbols())
    rho_syms = set(rho_p.symbols())
    B = poly_moment_matrix(U)
    H = []
    from fractions import Fraction
    s2 = Fraction(sigma2).limit_denominator(1000) if not isinstance(sigma2, Poly) else sigma2
    for j in range(U.shape[0]):
        h = rho_p
        for i in range(d):
            if B.p[j, i] == 1:
                h = h * Jv[i] * rho_p ** -1
            elif B.p[j, i] == 2:
                h = h * (Jv[i] ** 2 * rho_p ** -2 + s2)
        # order of a term = total |power| of every symbol except the density's (reference src/lib/feq.R:50-57): the
        # momentum may be given as a product such as d*u[0], whose density factor must not count
        count = [sy for sy in h.symbols() if sy not in rho_syms]
        H.append(h.truncate(count, order))
    if correction is not None:
        # added to the raw moments of order > 3, before any change of basis (reference src/lib/feq.R:59-63)
        sel = [j for j in range(U.shape[0]) if B.order[j] > 3]
        corr = list(correction)
        if len(sel) != len(corr):
            raise ValueError("Correction of wrong length in MRT_eq")
        for j, c in zip(sel, corr):
            H[j] = H[j] + c
    Req = PVec(H)
    Mat = B.mat
    ordr = B.order.copy()
    if mat is not None:
        T = np.linalg.solve(B.mat, np.asarray(mat, dtype=float))
        ordr = np.array([max(B.order[np.abs(T[:, j]) > 1e-10]) for j in range(T.shape[1])])
        Req = Req @ T
        Mat = np.asarray(mat, dtype=float)
    elif orthogonal:
        M = integer_orthogonalise(B.mat)
        Req = Req @ np.linalg.solve(B.mat, M)
        Mat = M
    feq = Req @ np.linalg.inv(Mat)
    return Equilibrium(U=U, Req=Req, mat=Mat, order=ordr, feq=feq, rho=rho_p, J=Jv)


def lattice_weights(U) -> np.ndarray:
    eq = mrt_eq(U, orthogonal=False)
    w = eq.feq.subst({"Jx": 0, "Jy": 0, "Jz": 0, "rho": 1})
    return np.array([float(x.const_value()) for x in w])


def weighted_cholesky_basis(U, scales: Sequence[float]) -> np.ndarray:
    """Moment matrix whose columns are mutually orthogonal w.r.t. the lattice weights,
    obtained by a reversed-order Cholesky factor of the weighted Gram inverse, with a per-
    moment normalisation.  (Used by the d2q9 family: reference
    ``models/flow/d2q9/Dynamics.c.Rt:8-23`` describes the construction.)"""
    B = poly_moment_matrix(np.asarray(U))
    wi = lattice_weights(U)
    Mi = np.linalg.inv(B.mat)
    W = Mi @ np.diag(1.0 / wi) @ Mi.T
    n = W.shape[0]
    r = np.arange(n)[::-1]
    Wr = W[np.ix_(r, r)]
    Lc = np.linalg.cholesky(Wr)          # Wr = L L^T ; upper factor R = L^T
    R = Lc.T
    H = R[np.ix_(r, r)]
    H = H * np.asarray(scales, dtype=float)[:, None]
    return B.mat @ H.T


# ------------------------------------------------------------------------------------------
# boundary closures
# ------------------------------------------------------------------------------------------

def bounce_index(U) -> np.ndarray:
    U = np.asarray(U, dtype=int)
    out = np.arange(U.shape[0])
    for i, u in enumerate(U):
        for j, v in enumerate(U):
            if np.all(u == -v):
                out[i] = j
    return out


def _pairs_by_group(densities, group: str, op: Callable[[np.ndarray], np.ndarray]):
    """Pairs (name_a, name_b, vec_a) of densities in the same group related by velocity map op."""
    groups: Dict[str, list] = {}
    for d in densities:
        groups.setdefault(d["group"], []).append(d)
    for g in sorted(groups):
        if group not in ("", None) and g != group and g not in (group if isinstance(group, (list, tuple)) else [group]):
            continue
        D = groups[g]
        for a in D:
            va = np.array([a["dx"], a["dy"], a["dz"]])
            vb = op(va)
            for b in D:
                if np.all(np.array([b["dx"], b["dy"], b["dz"]]) == vb) and a["name"] < b["name"]:
                    yield a["name"], b["name"], va


def full_bounce_back(densities, group: str = "", indent="\t") -> str:
    out = [f"{indent}real_t tmp;"]
    for a, b, _ in _pairs_by_group(densities, group, lambda v: -v):
        out += [f"{indent}tmp = {a};", f"{indent}{a} = {b};", f"{indent}{b} = tmp;"]
    return "\n".join(out) + "\n"


def symmetry(densities, direction: int, sign: int, group: str = "", indent="\t") -> str:
    """Mirror populations across the plane normal to axis `direction` (1-based)."""
    k = direction - 1

    def op(v):
        w = v.copy()
        w[k] = -w[k]
        return w
    out = []
    for a, b, va in _pairs_by_group(densities, group, op):
        # `a` sorts before `b`; decide the copy direction from the velocity of the pair
        # (the entry whose mirrored component points along `sign` receives the value)
        if sign * va[k] > 0:
            out.append(f"{indent}{a} = {b};")
        else:
            out.append(f"{indent}{b} = {a};")
    return "\n".join(out) + "\n"


def _solve_linear(eq: Poly, var: str) -> str:
    a, b = eq.split_linear(var)
    if a.is_zero():
        raise ValueError(f"equation does not contain {var}")
    if len(a.t) > 1:
        return f"( {to_c(-b)} ) / ( {to_c(a)} )"
    return to_c((-b) * a ** -1)


def zouhe(eq: Equilibrium, f: PVec, direction: int, sign: int, kind: str,
          P="Pressure", V="Velocity", V3: Optional[Sequence] = None,
          predefined: bool = False, indent="\t") -> str:
    """Non-equilibrium bounce-back closure for a face normal to axis `direction` (1-based);
    populations with sign*c_dir > 0 are unknown.  `kind` is "velocity" or "pressure"."""
    U = eq.U
    d = U.shape[1]
    k = direction - 1
    Js = ["Jx", "Jy", "Jz"][:d]
    Vp = Poly.lift(V)
    V3v = [Poly.const(0)] * d if V3 is None else [Poly.lift(v) for v in V3]
    if V3 is None:
        V3v[k] = Vp
    b = bounce_index(U)
    sel = sign * U[:, k] > 0
    feq = eq.feq
    fs = f.copy()
    for i in np.nonzero(sel)[0]:
        fs[i] = feq[i] + f[b[i]] - feq[b[i]]
    mass = fs.sum() - eq.rho
    mom = [psum(fs[i] * int(U[i, a]) for i in range(len(f)) if U[i, a] != 0) - Poly.sym(Js[a])
           for a in range(d)]
    out = []
    if not predefined:
        out.append(f"{indent}real_t Jx, Jy, Jz, rho;")
    rho_name = eq.rho.symbols()[0]
    if kind == "pressure":
        out.append(f"{indent}{rho_name} = {to_c(Poly.lift(P) * 3 + 1)};")
        out.append(f"{indent}{Js[k]} = {_solve_linear(mass, Js[k])};")
    elif kind == "velocity":
        nJ = eq.rho * Vp
        out.append(f"{indent}{rho_name} = {_solve_linear(mass.subst({Js[k]: nJ}), rho_name)};")
        out.append(f"{indent}{Js[k]} = {to_c(nJ)};")
    else:
        raise ValueError(kind)
    for a in range(d):
        if a != k:
            out.append(f"{indent}{Js[a]} = {_solve_linear(mom[a], Js[a])};")
    for a in range(d):
        if a != k and not V3v[a].is_zero():
            out.append(f"{indent}{Js[a]} = {to_c(Poly.sym(Js[a]) + eq.rho * V3v[a])};")
    for i in np.nonzero(sel)[0]:
        out.append(f"{indent}{to_c(f[i])} = {to_c(fs[i])};")
    return "\n".join(out) + "\n"


def zouhe_rewrite(U, f: PVec, n: Sequence[int], kind: str, rhs=None, indent="\t") -> str:
    """Zou-He closure written with an auxiliary momentum R: unknown populations (c.n < 0)
    get feq(R) + bounced non-equilibrium; R is solved from the prescribed density (pressure
    type) or velocity (velocity type)."""
    U = np.asarray(U, dtype=int)
    n = np.asarray(n, dtype=int)
    d = len(n)
    assert np.sum(n != 0) == 1 and np.sum(np.abs(n)) == 1, "normal must be a cartesian versor"
    k = int(np.nonzero(n)[0][0])
    if rhs is None:
        if kind == "velocity":
            rhs = PVec([Poly.sym("Velocity") * int(abs(x)) for x in n])
        else:
            rhs = Poly.sym("Pressure") * 3 + 1
    Rn = ["Rx", "Ry", "Rz"][:d]
    eq2 = mrt_eq(U, J=Rn, orthogonal=False)
    feq = eq2.feq
    sel = (U @ n) < 0
    b = bounce_index(U)
    fs = f.copy()
    for i in np.nonzero(sel)[0]:
        fs[i] = feq[i] + f[b[i]] - feq[b[i]]
    mom = [psum(fs[i] * int(U[i, a]) for i in range(len(f)) if U[i, a] != 0) for a in range(d)]
    if kind == "pressure":
        eqn = list(mom)
        eqn[k] = fs.sum()
        r = Poly.lift(rhs)
        rh = [r * int(abs(x)) for x in n]
    elif kind == "velocity":
        eqn = list(mom)
        tot = fs.sum()
        rv = rhs if isinstance(rhs, PVec) else PVec(rhs)
        rh = [rv[a] * tot for a in range(d)]
    else:
        raise ValueError(kind)
    eqn = [e - r for e, r in zip(eqn, rh)]
    out = [f"{indent}/* {kind}-type Zou-He closure, normal {tuple(int(x) for x in n)} */"]
    order = [k] + [a for a in range(d) if a != k]
    for a in order:
        out.append(f"{indent}real_t {Rn[a]} = {_solve_linear(eqn[a], Rn[a])};")
    for i in np.nonzero(sel)[0]:
        out.append(f"{indent}{to_c(f[i])} = {to_c(fs[i])};")
    return "\n".join(out) + "\n"


# ------------------------------------------------------------------------------------------
# d3q19 classic MRT (d'Humieres et al. 2002 basis)
# ------------------------------------------------------------------------------------------

def d3q19_velocities() -> np.ndarray:
    return np.array([
        [0, 0, 0],
        [1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1],
        [1, 1, 0], [-1, 1, 0], [1, -1, 0], [-1, -1, 0],
        [1, 0, 1], [-1, 0, 1], [1, 0, -1], [-1, 0, -1],
        [0, 1, 1], [0, -1, 1], [0, 1, -1], [0, -1, -1]], dtype=int)


def d3q19_mrt_matrix() -> np.ndarray:
    """M (19 moments x 19 velocities), built from the standard polynomial definitions
    rho, e, eps, jx, qx, jy, qy, jz, qz, 3pxx, 3pixx, pww, piww, pxy, pyz, pxz, mx, my, mz."""
    c = d3q19_velocities().astype(float)
    x, y, z = c[:, 0], c[:, 1], c[:, 2]
    c2 = x * x + y * y + z * z
    rows = [
        np.ones(19),
        19 * c2 - 30,
        (21 * c2 * c2 - 53 * c2 + 24) / 2,
        x, (5 * c2 - 9) * x,
        y, (5 * c2 - 9) * y,
        z, (5 * c2 - 9) * z,
        3 * x * x - c2, (3 * c2 - 5) * (3 * x * x - c2),
        y * y - z * z, (3 * c2 - 5) * (y * y - z * z),
        x * y, y * z, x * z,
        (y * y - z * z) * x, (z * z - x * x) * y, (x * x - y * y) * z,
    ]
    return np.array(rows)


# ------------------------------------------------------------------------------------------
# cumulant transforms for tensor-product velocity sets (d3q27)
# ------------------------------------------------------------------------------------------

def _name(prefix, a, b, c):
    return f"{prefix}{a}{b}{c}"


def raw_moment_sweeps(prefix: str = "f", forward: bool = True, indent="\t") -> str:
    """In-place separable transform between populations (index digit 0,1,2 = velocity 0,+1,-1)
    and raw moments (digit = power) along x, then y, then z (reverse order for the inverse)."""
    out = []
    axes = [0, 1, 2] if forward else [0, 1, 2]
    for ax in axes:
        for j, k in product(range(3), repeat=2):
            def nm(d):
                t = [0, 0, 0]
                others = [a for a in range(3) if a != ax]
                t[ax] = d
                t[others[0]] = j
                t[others[1]] = k
                return _name(prefix, *t)
            z, p, m = nm(0), nm(1), nm(2)
            if forward:
                # m0 = f0+f+ + f- ; m1 = f+ - f- ; m2 = f+ + f-
                out.append(f"{indent}{z} = {m} + {p} + {z}; {p} = -{m} + {p}; {m} = {p} + {m}*2.0;")
            else:
                # f0 = m0 - m2 ; f+ = (m2+m1)/2 ; f- = m2 - f+
                out.append(f"{indent}{z} = -{m} + {z}; {p} = ( {m} + {p} )/2.0; {m} = {m} - {p};")
    return "\n".join(out) + "\n"


def _multi_binom(n, k):
    from math import comb
    r = 1
    for a, b in zip(n, k):
        r *= comb(a, b)
    return r


def cumulant_from_moments(mprefix="f", cprefix="c", indent="\t") -> str:
    """c_alpha from raw moments m (stored in the population variables), using
       m_alpha = sum_{beta<=alpha, beta_j>=1} C(alpha-e_j, beta-e_j) c_beta m_{alpha-beta}
    normalised by m_000 (so c_000 is not formed and c_100 = m_100/m_000 ...)."""
    lines = []
    m0 = _name(mprefix, 0, 0, 0)
    for c, b, a in product(range(3), repeat=3):      # a fastest
        al = (a, b, c)
        if al == (0, 0, 0):
            continue
        j = next(i for i in range(3) if al[i] > 0)
        ej = tuple(1 if i == j else 0 for i in range(3))
        terms: Dict[int, list] = {}
        for be in product(*[range(x + 1) for x in al]):
            if be[j] < 1 or be == al:
                continue
            coef = _multi_binom(tuple(x - e for x, e in zip(al, ej)), tuple(x - e for x, e in zip(be, ej)))
            rest = tuple(x - y for x, y in zip(al, be))
            terms.setdefault(coef, []).append(f"{_name(cprefix, *be)}*{_name(mprefix, *rest)}")
        expr = _name(mprefix, *al)
        for coef in sorted(terms):
            s = " - ".join(terms[coef])
            expr += f" - {s}" if coef == 1 else f" - ( {' + '.join(terms[coef])} )*{coef}.0"
        lines.append(f"{indent}{_name(cprefix, *al)} = ( {expr} )/{m0};")
    return "\n".join(lines) + "\n"


def moments_from_cumulant(mprefix="f", cprefix="c", indent="\t") -> str:
    lines = []
    m0 = _name(mprefix, 0, 0, 0)
    for c, b, a in product(range(3), repeat=3):
        al = (a, b, c)
        if al == (0, 0, 0):
            continue
        j = next(i for i in range(3) if al[i] > 0)
        ej = tuple(1 if i == j else 0 for i in range(3))
        terms: Dict[int, list] = {}
        for be in product(*[range(x + 1) for x in al]):
            if be[j] < 1:
                continue
            coef = _multi_binom(tuple(x - e for x, e in zip(al, ej)), tuple(x - e for x, e in zip(be, ej)))
            rest = tuple(x - y for x, y in zip(al, be))
            terms.setdefault(coef, []).append(f"{_name(cprefix, *be)}*{_name(mprefix, *rest)}")
        parts = []
        for coef in sorted(terms):
            s = " + ".join(terms[coef])
            parts.append(s if coef == 1 else f"( {s} )*{coef}.0")
        lines.append(f"{indent}{_name(mprefix, *al)} = {' + '.join(parts)};")
    return "\n".join(lines) + "\n"


# ------------------------------------------------------------------------------------------
# central-moment sweeps (B200: the cumulant kernel is FP32-issue bound, so the transform is
# written with the minimum number of dependent FMAs: 6 per triple forward, 7 backward)
# ------------------------------------------------------------------------------------------

def _triple_names(prefix, ax, j, k):
    def nm(d):
        t = [0, 0, 0]
        others = [a for a in range(3) if a != ax]
        t[ax] = d
        t[others[0]] = j
        t[others[1]] = k
        return _name(prefix, *t)
    return nm(0), nm(1), nm(2)


def central_moment_forward(prefix: str, u: Sequence[str], indent="\t") -> str:
    """populations (digit 0,1,2 = velocity 0,+1,-1) -> central moments about u (digit = power), in place."""
    out = [f"{indent}real_t cm_a, cm_b;"]
    for ax in range(3):
        for j, k in product(range(3), repeat=2):
            z, p, m = _triple_names(prefix, ax, j, k)
            out.append(f"{indent}cm_a = {p} + {m}; cm_b = {p} - {m}; {z} = {z} + cm_a; "
                       f"{p} = cm_b - {u[ax]}*{z}; {m} = cm_a - {u[ax]}*(cm_b + {p});")
    return "\n".join(out) + "\n"


def central_moment_backward(prefix: str, u: Sequence[str], indent="\t") -> str:
    """central moments about u -> populations, in place (inverse of central_moment_forward)."""
    out = [f"{indent}real_t cr_1, cr_2;"]
    for ax in range(3):
        for j, k in product(range(3), repeat=2):
            z, p, m = _triple_names(prefix, ax, j, k)
            out.append(f"{indent}cr_1 = {p} + {u[ax]}*{z}; cr_2 = {m} + {u[ax]}*({p} + cr_1); {z} = {z} - cr_2; "
                       f"{p} = (cr_2 + cr_1)*0.5; {m} = cr_2 - {p};")
    return "\n".join(out) + "\n"


def central_moment_forward_shared(prefix: str, rho: str, j: Sequence[str], u: Sequence[str], irho: str, indent="\t",
                                  after_velocity: str = "") -> str:
    """Forward central-moment transform that also produces density and momentum from the partial sums of
    its own first (x) pass: rho and J cost 8+8+5+5 additions instead of 26+3*17.
    `after_velocity` is code placed right after rho, J and u exist and before the sweeps need registers for their
    temporaries (the place for Globals that only need density and velocity)."""
    out = [f"{indent}real_t cm_a, cm_b;"]
    ax = 0
    for jj, k in product(range(3), repeat=2):
        z, p, m = _triple_names(prefix, ax, jj, k)
        out.append(f"{indent}cm_a = {p} + {m}; {p} = {p} - {m}; {m} = cm_a; {z} = {z} + cm_a;")
    def nm(a, b, c):
        return _name(prefix, a, b, c)
    zs = [nm(0, b, c) for c in range(3) for b in range(3)]
    out.append(f"{indent}real_t {rho} = ((({zs[0]} + {zs[1]}) + ({zs[2]} + {zs[3]})) + (({zs[4]} + {zs[5]}) + ({zs[6]} + {zs[7]}))) + {zs[8]};")
    bs = [nm(1, b, c) for c in range(3) for b in range(3)]
    out.append(f"{indent}real_t {j[0]} = ((({bs[0]} + {bs[1]}) + ({bs[2]} + {bs[3]})) + (({bs[4]} + {bs[5]}) + ({bs[6]} + {bs[7]}))) + {bs[8]};")
    out.append(f"{indent}real_t {j[1]} = (({nm(0,1,0)} - {nm(0,2,0)}) + ({nm(0,1,1)} - {nm(0,2,1)})) + ({nm(0,1,2)} - {nm(0,2,2)});")
    out.append(f"{indent}real_t {j[2]} = (({nm(0,0,1)} - {nm(0,0,2)}) + ({nm(0,1,1)} - {nm(0,1,2)})) + ({nm(0,2,1)} - {nm(0,2,2)});")
    out.append(f"{indent}real_t {irho} = 1.0/{rho};")
    out.append(f"{indent}{u[0]} = {j[0]}*{irho}; {u[1]} = {j[1]}*{irho}; {u[2]} = {j[2]}*{irho};")
    if after_velocity:
        out.append(after_velocity.rstrip("\n"))
    for jj, k in product(range(3), repeat=2):
        z, p, m = _triple_names(prefix, ax, jj, k)
        out.append(f"{indent}cm_b = {p}; {p} = cm_b - {u[0]}*{z}; {m} = {m} - {u[0]}*(cm_b + {p});")
    for ax in (1, 2):
        for jj, k in product(range(3), repeat=2):
            z, p, m = _triple_names(prefix, ax, jj, k)
            out.append(f"{indent}cm_a = {p} + {m}; cm_b = {p} - {m}; {z} = {z} + cm_a; "
                       f"{p} = cm_b - {u[ax]}*{z}; {m} = cm_a - {u[ax]}*(cm_b + {p});")
    return "\n".join(out) + "\n"


# ------------------------------------------------------------------------------------------
# moment transforms that exploit the inversion symmetry of the velocity set
# ------------------------------------------------------------------------------------------

def _group_terms(pairs, tol=1e-12):
    """[(coef, name)] -> C text  c1*(a + b) + c2*(c) ... with +-1 coefficients folded into signs."""
    groups: Dict[float, list] = {}
    for c, n in pairs:
        if abs(c) < tol:
            continue
        key = None
        for k in groups:
            if abs(abs(k) - abs(c)) < tol * max(1.0, abs(c)):
                key = k
                break
        if key is None:
            key = abs(c)
            groups[key] = []
        groups[key].append(("-" if c < 0 else "+", n))
    parts = []
    for k in sorted(groups, key=lambda v: -v):
        mem = groups[k]
        body = ""
        for i, (sg, n) in enumerate(mem):
            body += (("-" if sg == "-" else "") if i == 0 else (" - " if sg == "-" else " + ")) + n
        one = abs(k - 1.0) < tol
        if one:
            parts.append(body if len(mem) == 1 else f"({body})")
        else:
            from .poly import _fmt_num, _clean
            parts.append(f"({body})*{_fmt_num(_clean(k))}")
    if not parts:
        return "0.0"
    s = parts[0]
    for p in parts[1:]:
        s += (" - " + p[1:]) if p.startswith("-") and not p.startswith("-(") and not p.startswith("(-") else " + " + p
    return s


def sym_moments(f: PVec, M: np.ndarray, U: np.ndarray, out, prefix="sy", indent="\t") -> str:
    """out_j = sum_k f_k M[k,j], computed from pair sums s_k = f_k + f_opp(k) and differences d_k = f_k - f_opp(k):
    even moments only need the sums, odd moments only the differences, which halves the multiply-adds of a dense
    product; equal coefficients are grouped so each distinct value is multiplied once."""
    b = bounce_index(U)
    Q = len(f)
    names = [to_c(x) for x in f]
    lines = []
    rep = {}
    for k in range(Q):
        if b[k] == k:
            rep[k] = ("self", names[k])
        elif k < b[k]:
            s, d = f"{prefix}s{k}", f"{prefix}d{k}"
            lines.append(f"{indent}const real_t {s} = {names[k]} + {names[b[k]]}, {d} = {names[k]} - {names[b[k]]};")
            rep[k] = ("pair", s, d)
    outs = [to_c(o) if not isinstance(o, str) else o for o in out]
    for j in range(M.shape[1]):
        terms = []
        for k in range(Q):
            if b[k] == k:
                terms.append((float(M[k, j]), rep[k][1]))
            elif k < b[k]:
                e = 0.5 * (M[k, j] + M[b[k], j])
                o = 0.5 * (M[k, j] - M[b[k], j])
                terms.append((float(e), rep[k][1]))
                terms.append((float(o), rep[k][2]))
        lines.append(f"{indent}{outs[j]} = {_group_terms(terms)};")
    return "\n".join(lines) + "\n"


def sym_populations(R, B: np.ndarray, U: np.ndarray, f: PVec, prefix="sy", indent="\t") -> str:
    """f_k = sum_j R_j B[j,k] for a velocity set with inversion symmetry: the even and odd parts of each pair are
    formed once, f_k = E + O, f_opp(k) = E - O."""
    b = bounce_index(U)
    Q = len(f)
    names = [to_c(x) for x in f]
    Rn = [to_c(r) if not isinstance(r, str) else r for r in R]
    lines = []
    for k in range(Q):
        if b[k] == k:
            terms = [(float(B[j, k]), Rn[j]) for j in range(B.shape[0])]
            lines.append(f"{indent}{names[k]} = {_group_terms(terms)};")
        elif k < b[k]:
            te = [(float(0.5 * (B[j, k] + B[j, b[k]])), Rn[j]) for j in range(B.shape[0])]
            to = [(float(0.5 * (B[j, k] - B[j, b[k]])), Rn[j]) for j in range(B.shape[0])]
            E, O = f"{prefix}E{k}", f"{prefix}O{k}"
            lines.append(f"{indent}const real_t {E} = {_group_terms(te)};")
            lines.append(f"{indent}const real_t {O} = {_group_terms(to)};")
            lines.append(f"{indent}{names[k]} = {E} + {O}; {names[b[k]]} = {E} - {O};")
    return "\n".join(lines) + "\n"
