"""Native counterparts of the generator library that model files pull in with ``source("lib/...")``.

The reference ships R implementations (``src/lib/feq.R``, ``boundary.R``, ``lattice.R``, ``d3q19.R``, ``FMT.R``,
``utils.R``) built on the external polyAlgebra/gvector packages.  Here the same R-level names are bound to this
framework's own symbolic layer (:mod:`tclb_b200.gen.lbmlib`, :mod:`tclb_b200.gen.poly`), so a model directory needs
nothing but its own files.  Each ``install_*`` function is what ``source("lib/<file>")`` executes.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Any, Dict, List, Sequence

import numpy as np

from . import lbmlib
from .poly import Poly, PVec, c_assign, to_c
from .rlang import NULL, DataFrame, RError, RList, scalar, truthy, vec


def _args(names: Sequence[str], a: list, k: Dict[str, Any], defaults: Dict[str, Any] = None) -> Dict[str, Any]:
    """R-style argument matching for a native function: exact names, unique prefixes, then positions."""
    out: Dict[str, Any] = {}
    kw = dict(k)
    for key in list(kw):
        if key in names:
            out[key] = kw.pop(key)
    for key in list(kw):
        cand = [n for n in names if n.startswith(key) and n not in out]
        if len(cand) == 1:
            out[cand[0]] = kw.pop(key)
    if kw:
        raise RError(f"unused argument(s): {', '.join(kw)}")
    free = [n for n in names if n not in out]
    for v in a:
        if not free:
            raise RError("too many arguments")
        out[free.pop(0)] = v
    for n, d in (defaults or {}).items():
        out.setdefault(n, d)
    return out


def _mat(x) -> np.ndarray:
    if isinstance(x, DataFrame):
        return np.array([[float(v) for v in col] for col in x.cols.values()]).T
    m = np.asarray(vec(x), dtype=float)
    return m


def _pvec(x) -> PVec:
    if isinstance(x, PVec):
        return x
    if isinstance(x, Poly):
        return PVec([x])
    out = []
    for v in vec(x).ravel(order="F").tolist():
        out.append(Poly.sym(v) if isinstance(v, str) else Poly.const(float(v)))
    return PVec(out)


def _densities(I) -> List[dict]:
    m = getattr(I.globals, "model", None)
    if m is not None:
        return [d for d in m.density_dicts() if not d.get("adjoint")]
    # no ModelDef behind the interpreter (the reference's own conf.R was sourced): read its Density table
    D = I.globals.get("Density")
    c = D.cols
    return [dict(name=c["name"][i], dx=int(c["dx"][i]), dy=int(c["dy"][i]), dz=int(c["dz"][i]), group=c["group"][i]) for i in range(D.nrow())]


# ---------------------------------------------------------------------------------------------- lattice.R
def d2q9_matrix(U: np.ndarray) -> np.ndarray:
    """Lallemand-Luo moment set, rows in the order (1, cx, cy, e, eps, qx, qy, pxx, pxy); returned as columns."""
    cx, cy = U[:, 0].astype(float), U[:, 1].astype(float)
    c2 = cx * cx + cy * cy
    rows = [np.ones_like(cx), cx, cy, -4 + 3 * c2, 4 - 10.5 * c2 + 4.5 * c2 * c2, (-5 + 3 * c2) * cx, (-5 + 3 * c2) * cy,
            cx * cx - cy * cy, cx * cy]
    return np.array(rows).T


def d3q7_matrix(U: np.ndarray) -> np.ndarray:
    cx, cy, cz = (U[:, i].astype(float) for i in range(3))
    c2 = cx * cx + cy * cy + cz * cz
    return np.array([np.ones_like(cx), cx, cy, cz, cx * cx - cy * cy, cx * cx + cy * cy - 2 * cz * cz, -6 + 7 * c2]).T


def velocity_sets() -> Dict[str, np.ndarray]:
    d2q9 = np.zeros((9, 3), dtype=int)
    d2q9[:, :2] = lbmlib.d2q9_velocities()
    d3q7 = np.array([[0, 0, 0], [1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]])
    d3q19 = lbmlib.d3q19_velocities()
    # d3q27 in the order used by the phase-field models: rest, 6 axis, 8 corners, 12 edges (xy, xz, yz planes)
    d3q27 = [[0, 0, 0], [1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]
    d3q27 += [[sx, sy, sz] for sz in (1, -1) for sy in (1, -1) for sx in (1, -1)]
    d3q27 += [[sx, sy, 0] for sy in (1, -1) for sx in (1, -1)]
    d3q27 += [[sx, 0, sz] for sz in (1, -1) for sx in (1, -1)]
    d3q27 += [[0, sy, sz] for sz in (1, -1) for sy in (1, -1)]
    return dict(d2q9=d2q9, d3q7=d3q7, d3q19=np.asarray(d3q19), d3q27=np.array(d3q27))


def known_matrix(U: np.ndarray):
    """the canonical MRT matrix that lattice.R attaches to its velocity sets (attribute MAT), if U is one of them"""
    U = np.asarray(U)
    sets = velocity_sets()
    for name, fn in (("d2q9", lambda: d2q9_matrix(sets["d2q9"])), ("d3q7", lambda: d3q7_matrix(sets["d3q7"])),
                     ("d3q19", _d3q19_matrix)):
        S = sets[name]
        if U.shape == S.shape and np.array_equal(U.astype(int), S):
            return fn()
    return None


def install_lattice(I, env):
    for name, U in velocity_sets().items():
        env.set(name, U.astype(float))
    return NULL


# ---------------------------------------------------------------------------------------------- feq.R
def _eq_to_rlist(eq: lbmlib.Equilibrium, extra: Dict[str, Any] = None) -> RList:
    r = RList()
    r["Req"] = eq.Req
    r["mat"] = np.asarray(eq.mat, dtype=float)
    r["p"] = NULL
    r["order"] = np.asarray(eq.order, dtype=float)
    r["U"] = np.asarray(eq.U, dtype=float)
    U = np.asarray(eq.U, dtype=float)
    d = U.shape[1]
    D2 = np.zeros((U.shape[0], d, d))
    for i in range(d):
        for j in range(d):
            D2[:, i, j] = U[:, i] * U[:, j]
    r["D2"] = D2
    for kk, v in (extra or {}).items():
        r[kk] = v
    r["feq"] = eq.feq
    return r


def install_feq(I, env):
    def mrt_polymatrix(I_, a, k):
        U = _mat(a[0]).astype(int)
        B = lbmlib.poly_moment_matrix(U)
        p = np.where(U < 0, 2, U)
        can = np.argsort(p.sum(axis=1), kind="stable") + 1
        r = RList()
        r["order"] = B.order.astype(float)
        r["mat"] = B.mat
        r["p"] = B.p.astype(float)
        r["canonical"] = can.astype(float)
        return r

    def mrt_eq(I_, a, k):
        g = _args(["U", "rho", "J", "sigma2", "order", "ortogonal", "mat", "correction"], a, k)
        U = _mat(g["U"]).astype(int)
        rho = _pvec(g["rho"])[0] if "rho" in g else Poly.sym("rho")
        J = _pvec(g["J"]) if "J" in g else PVec(["Jx", "Jy", "Jz"])
        sigma2 = float(scalar(g["sigma2"])) if "sigma2" in g and not isinstance(g["sigma2"], (Poly, PVec)) else g.get("sigma2", 1.0 / 3.0)
        if isinstance(sigma2, PVec):
            sigma2 = sigma2[0]
        order = int(scalar(g["order"])) if "order" in g else 2
        orth = truthy(g["ortogonal"]) if "ortogonal" in g else True
        mat = _mat(g["mat"]) if "mat" in g and g["mat"] is not NULL else known_matrix(U)
        try:
            eq = lbmlib.mrt_eq(U, rho=rho, J=J, sigma2=sigma2, order=order, orthogonal=orth, mat=mat,
                               correction=_pvec(g["correction"]) if "correction" in g else None)
        except ValueError as e:
            raise RError(str(e))
        extra = {}
        if mat is None and not orth:
            B = lbmlib.poly_moment_matrix(U)
            extra["p"] = B.p.astype(float)
            p = np.where(U < 0, 2, U)
            extra["canonical"] = (np.argsort(p.sum(axis=1), kind="stable") + 1).astype(float)
        r = _eq_to_rlist(eq, None)
        for kk, v in extra.items():
            r[kk] = v
        return r

    def wmrt_mat(I_, a, k):
        """WMRT_mat(U) (src/lib/feq.R:90-113): the polynomial moment matrix recombined so that the high-order equilibrium
        becomes "triangular" in the monomials of (rho, J): M %*% solve(R), R = unit-diagonal upper factor of the QR
        decomposition of the table of coefficients of the equilibrium moments."""
        U = _mat(a[0]).astype(int)
        try:
            eq = lbmlib.mrt_eq(U, rho=Poly.sym("rho"), J=PVec(["Jx", "Jy", "Jz"]), sigma2=1.0 / 3.0, order=300, orthogonal=False, mat=None)
        except ValueError as e:
            raise RError(str(e))
        M = np.asarray(eq.mat, dtype=float)
        monos = []
        for p in eq.Req:
            for mono in p.t:
                if mono not in monos:
                    monos.append(mono)
        n = len(eq.Req)
        A = np.array([[float(p.t.get(mono, 0.0)) for p in eq.Req] for mono in monos], dtype=float)
        R = np.eye(n)
        r = np.linalg.qr(A, mode="r")
        R[:r.shape[0], :] = r[:n, :]
        R = R / np.diag(R)[:, None]
        return M @ np.linalg.inv(R)

    env.set("WMRT_mat", wmrt_mat)
    env.set("MRT_polyMatrix", mrt_polymatrix)
    env.set("MRT_integerOrtogonal", lambda I_, a, k: lbmlib.integer_orthogonalise(_mat(a[0])))
    env.set("MRT_eq", mrt_eq)
    env.set("MRT_feq", lambda I_, a, k: mrt_eq(I_, a, k)["feq"])

    return NULL


# ---------------------------------------------------------------------------------------------- boundary.R
def _emit(I, text: str):
    I.out.append(text)


def _c_pull(I, eq: Poly, var: str):
    _emit(I, f"{var} = {lbmlib._solve_linear(eq, var)};\n")


def install_boundary(I, env):
    def bounce(I_, a, k):
        return (lbmlib.bounce_index(_mat(a[0]).astype(int)) + 1).astype(float)

    def group_arg(g):
        if g is None or g is NULL:
            return ""
        v = [x for x in vec(g).tolist()]
        return v[0] if len(v) == 1 else v

    def full_bounce_back(I_, a, k):
        g = _args(["group"], a, k)
        _emit(I_, lbmlib.full_bounce_back(_densities(I_), group_arg(g.get("group")), indent=""))
        return NULL

    def symmetry(I_, a, k):
        g = _args(["direction", "sign", "group"], a, k)
        _emit(I_, lbmlib.symmetry(_densities(I_), int(scalar(g["direction"])), int(scalar(g["sign"])), group_arg(g.get("group")), indent=""))
        return NULL

    def c_pull(I_, a, k):
        W = _pvec(a[0])
        _c_pull(I_, W[0], scalar(a[1]))
        return NULL

    def eq_from(x) -> lbmlib.Equilibrium:
        if not isinstance(x, RList):
            raise RError("an equilibrium object (result of MRT_eq) is required")
        return lbmlib.Equilibrium(U=np.asarray(x["U"]).astype(int), Req=x["Req"], mat=np.asarray(x["mat"], dtype=float),
                                  order=np.asarray(x["order"]).astype(int), feq=x["feq"], rho=Poly.sym("rho"), J=PVec(["Jx", "Jy", "Jz"]))

    def zouhe(I_, a, k):
        g = _args(["EQ", "direction", "sign", "type", "group", "P", "V", "V3", "predefined"], a, k)
        eq = eq_from(g["EQ"])
        f = _pvec(g["group"]) if "group" in g else _pvec(env.get("f"))
        P = _pvec(g["P"])[0] if "P" in g else Poly.sym("Pressure")
        V = _pvec(g["V"])[0] if "V" in g else Poly.sym("Velocity")
        V3 = list(_pvec(g["V3"])) if "V3" in g else None
        pre = scalar(g["predefined"]) if "predefined" in g else "false"
        pre = (pre is True) or (isinstance(pre, str) and pre.lower() == "true") or (isinstance(pre, (bool, np.bool_)) and bool(pre))
        _emit(I_, lbmlib.zouhe(eq, f, int(scalar(g["direction"])), int(scalar(g["sign"])), scalar(g["type"]), P=P, V=V, V3=V3,
                               predefined=pre, indent=""))
        return NULL

    def zouhe_rewrite(I_, a, k):
        g = _args(["EQ", "f", "n", "type", "rhs"], a, k)
        EQ = g["EQ"]
        U = np.asarray(EQ["U"]).astype(int) if isinstance(EQ, RList) else _mat(EQ).astype(int)
        n = [int(v) for v in vec(g["n"]).tolist()]
        kind = scalar(g["type"]) if "type" in g else "velocity"
        rhs = g.get("rhs")
        if rhs is not None:
            rhs = _pvec(rhs)
            rhs = rhs if len(rhs) > 1 else rhs[0]
        _emit(I_, lbmlib.zouhe_rewrite(U, _pvec(g["f"]), n, kind, rhs, indent=""))
        return NULL

    def zouhe_new(I_, a, k):
        """ZouHeNew(EQ, f, direction, sign, order, group=f, known="rho", mom): the unknown populations become the
        bounced non-equilibrium closure, the moments up to `order` are matched and the equations are solved one
        unknown at a time (reference src/lib/boundary.R:118-150)."""
        g = _args(["EQ", "f", "direction", "sign", "order", "group", "known", "mom"], a, k)
        eq = eq_from(g["EQ"])
        f = _pvec(g["f"])
        fs = _pvec(g["group"]) if "group" in g else f
        direction, sign, order = int(scalar(g["direction"])), int(scalar(g["sign"])), int(scalar(g["order"]))
        U = np.asarray(eq.U).astype(int)
        bounce = lbmlib.bounce_index(U)
        sel = sign * U[:, direction - 1] > 0
        feq = eq.feq
        new = fs.copy()
        for i in np.nonzero(sel)[0]:
            new[int(i)] = feq[int(i)] + (fs[int(bounce[i])] - feq[int(bounce[i])])
        Rs = new @ np.asarray(eq.mat, dtype=float)
        low = [i for i in range(len(eq.order)) if eq.order[i] <= order]
        mom = _pvec(g["mom"]) if "mom" in g else None
        eqs = [(mom[j] if mom is not None else eq.Req[i]) - Rs[i] for j, i in enumerate(low)]
        known = set(str(v) for v in vec(g["known"]).tolist()) if "known" in g else {"rho"}
        known |= {to_c(f[int(i)]) for i in np.nonzero(~sel)[0]}
        needed = []
        for p in new:
            for sy in p.symbols():
                if sy not in known and sy not in needed:
                    needed.append(sy)
        while needed:
            pick = None
            for e_i, e in enumerate(eqs):
                present = [sy for sy in needed if sy in e.symbols()]
                if len(present) == 1:
                    pick = (e_i, present[0])
                    break
            if pick is None:
                raise RError("Cannot solve the boundary problem in ZouHe")
            e_i, sy = pick
            _c_pull(I_, eqs[e_i], sy)
            needed.remove(sy)
            eqs.pop(e_i)
        _emit(I_, c_assign([f[int(i)] for i in np.nonzero(sel)[0]], [new[int(i)] for i in np.nonzero(sel)[0]], indent="", optimize=False))
        return NULL

    env.set("Bounce", bounce)
    env.set("FullBounceBack", full_bounce_back)
    env.set("Symmetry", symmetry)
    env.set("C_pull", c_pull)
    env.set("ZouHe", zouhe)
    env.set("ZouHeRewrite", zouhe_rewrite)
    env.set("ZouHeNew", zouhe_new)
    return NULL


# ---------------------------------------------------------------------------------------------- d3q19.R
def _d3q19_matrix() -> np.ndarray:
    """d'Humieres moment matrix with one COLUMN per moment (column 1 = ones, columns 4/6/8 = velocity components)"""
    M = np.asarray(lbmlib.d3q19_mrt_matrix(), dtype=float)
    return M if np.all(M[:, 0] == 1) else M.T


def install_d3q19(I, env):
    M = _d3q19_matrix()
    env.set("d3q19_MRTMAT", M)

    def d3q19_mrt(I_, a, k):
        g = _args(["rho", "J", "R", "group"], a, k)
        rho = _pvec(g["rho"])[0] if "rho" in g else Poly.sym("rho")
        J = _pvec(g["J"]) if "J" in g else PVec(["Jx", "Jy", "Jz"])
        U = M[:, [3, 5, 7]].astype(int)
        eq = lbmlib.mrt_eq(U, rho=rho, J=J, order=2, mat=M)
        selU = [3, 5, 7]
        selR = [i for i in range(19) if i not in [0] + selU]
        Rs = PVec([rho] * 19)
        for i, j in zip(selU, J):
            Rs[i] = j
        R = g.get("R", vec("R"))
        if not isinstance(R, PVec):
            names = [str(v) for v in vec(R).tolist()]
            R = PVec([f"{names[0]}{i}" for i in range(1, 20)]) if len(names) == 1 else PVec(names)
        if len(R) == len(selR):
            for i, r in zip(selR, R):
                Rs[i] = r
        elif len(R) == 19:
            for i in selR:
                Rs[i] = R[i]
        else:
            raise RError("Bad length of R in d3q19_MRT")
        out = RList()
        out["MAT"] = M
        out["Req"] = eq.Req
        out["R"] = Rs
        if "group" in g:
            grp = scalar(g["group"])
            out["f"] = PVec([d["name"] for d in _densities(I_) if d["group"] == grp])
        else:
            out["f"] = NULL
        out["rho"] = PVec([rho])
        out["J"] = J
        out["U"] = U.astype(float)
        out["selR"] = np.array([i + 1 for i in selR], dtype=float)
        return out
    env.set("d3q19_MRT", d3q19_mrt)
    return NULL


# ---------------------------------------------------------------------------------------------- FMT.R
def install_fmt(I, env):
    m = np.array([[1, 1, 1], [0, 1, -1], [0, 1, 1]], dtype=float)

    def lu_nopivot(A):
        n = A.shape[0]
        L, U = np.eye(n), A.astype(float).copy()
        for c in range(n):
            for r in range(c + 1, n):
                L[r, c] = U[r, c] / U[c, c]
                U[r] -= L[r, c] * U[c]
        return L, U
    Lf, Uf = lu_nopivot(m)
    fmt_mat = np.eye(3) - np.linalg.inv(Lf) + Uf
    Li, Ui = lu_nopivot(np.linalg.inv(m))
    fmt_inv = np.eye(3) - np.linalg.inv(Li) + Ui
    env.set("FMT_mat", fmt_mat)
    env.set("FMT_inv", fmt_inv)

    def fmt(I_, a, k):
        g = _args(["I", "M", "sel", "inverse"], a, k)
        idx = _mat(g["I"]).astype(int)
        Mv = _pvec(g["M"])
        inv = truthy(g["inverse"]) if "inverse" in g else False
        mat = fmt_inv if inv else fmt_mat
        sel = np.asarray(vec(g["sel"]), dtype=bool) if "sel" in g else np.ones(idx.shape[0], dtype=bool)
        for s in range(idx.shape[1]):
            others = [c for c in range(idx.shape[1]) if c != s]
            groups: "OrderedDict[tuple, list]" = OrderedDict()
            for r in range(idx.shape[0]):
                groups.setdefault(tuple(idx[r, others][::-1]), []).append(r)
            for key in sorted(groups):
                rows = groups[key]
                if all(sel[r] for r in rows):
                    lhs = [Mv[r] for r in rows]
                    rhs = [sum((Mv[rows[j]] * float(mat[i, j]) for j in range(len(rows))), Poly()) for i in range(len(rows))]
                    _emit(I_, "".join(f"{to_c(l)} = {to_c(r)}; " for l, r in zip(lhs, rhs)))
                    _emit(I_, "\n")
        return NULL
    env.set("FMT", fmt)
    nums = RList()
    nums["1"] = np.array([1, 2, 3], dtype=float)
    nums["2"] = np.array([1, 2, 4, 3, 6, 7, 5, 9, 8], dtype=float)
    nums["3"] = np.array([1, 2, 3, 4, 8, 9, 5, 10, 11, 6, 12, 13, 16, 20, 21, 17, 22, 23, 7, 14, 15, 18, 24, 25, 19, 26, 27], dtype=float)
    env.set("FMT_num", nums)
    return NULL


# ---------------------------------------------------------------------------------------------- utils.R
def install_utils(I, env):
    def rows(I_, a, k):
        x = a[0]
        if isinstance(x, DataFrame):
            return RList.make([x.row(i) for i in range(x.nrow())])
        return x

    state = {"mark": False}

    def ifdef(I_, a, k):
        g = _args(["val", "tag"], a, k)
        val = truthy(g["val"]) if "val" in g else False
        tag = scalar(g["tag"]) if "tag" in g else "ADJOINT"
        if not state["mark"] and val:
            _emit(I_, f"\n#ifdef {tag} \n")
        if state["mark"] and not val:
            _emit(I_, f"\n#endif // {tag} \n")
        state["mark"] = val
        return NULL

    def c_table_decl(I_, a, k):
        import re
        g = _args(["d", "sizes"], a, k)
        names = [to_c(p) for p in g["d"]] if isinstance(g["d"], PVec) else [str(v) for v in vec(g["d"]).tolist()]
        sizes = truthy(g["sizes"]) if "sizes" in g else True
        arrays: "OrderedDict[str, int]" = OrderedDict()
        plain = []
        for n in names:
            mm = re.match(r"^\s*([^\[]*?)\s*\[\s*([^\] ]*)\s*\]", n)
            if mm:
                arrays[mm.group(1)] = max(arrays.get(mm.group(1), 0), int(mm.group(2)) + 1)
            else:
                plain.append(n)
        out = [f"{b}[{n}]" if sizes else b for b, n in sorted(arrays.items())] + plain
        return vec(out)
    env.set("rows", rows)
    env.set("ifdef", ifdef)
    env.set("c_table_decl", c_table_decl)
    return NULL


# ---------------------------------------------------------------------------------------------- cumulant.R
def install_cumulant(I, env):
    """FCT_eq(I, M, K) (src/lib/cumulant.R:7-44): moments expressed through cumulants by the recursion
    m_p = sum over subsets S of p \\ {s} of m_S * k_(p \\ S), for every multi-index row of I (after the first); of the
    len(p) equivalent forms (one per choice of the distinguished element s) the one with the fewest factors is kept."""
    import itertools
    axes = ["X", "Y", "Z"]

    def ord_(p):                      # 1-based position of a multiset of axes in the 3x3x3 tables
        return sum(p.count(v) * 3 ** k for k, v in enumerate(axes)) + 1

    def fct_eq(I_, a, k):
        g = _args(["I", "M", "K"], a, k)
        idx = _mat(g["I"]).astype(int)
        Mv, Kv = _pvec(g["M"]), _pvec(g["K"])
        out = Mv.copy()
        for r in range(1, idx.shape[0]):
            p = [axes[c] for c in range(idx.shape[1]) for _ in range(int(idx[r, c]))]
            n = len(p)
            subsets = [tuple(bool((j >> b) & 1) ^ True for b in range(n)) for j in range(2 ** n)]   # expand.grid(c(TRUE,FALSE)...): first column fastest
            cands = []
            for s_ in range(n):
                tot = Poly()
                for sel in subsets:
                    if sel[s_]:
                        continue
                    inside = [p[i] for i in range(n) if sel[i]]
                    outside = [p[i] for i in range(n) if not sel[i]]
                    tot = tot + Mv[ord_(inside) - 1] * Kv[ord_(outside) - 1]
                cands.append(tot)
            cost = [len(to_c(c).split("*")) for c in cands]
            out[ord_(p) - 1] = cands[int(np.argmin(cost))]
        return out

    env.set("FCT_eq", fct_eq)
    env.set("all.subsets", lambda I_, a, k: np.array([[bool((j >> b) & 1) ^ True for b in range(int(scalar(a[0])))] for j in range(2 ** int(scalar(a[0])))], dtype=bool))
    return NULL


HOOKS = {
    "lib/cumulant.R": install_cumulant,
    "lib/lattice.R": install_lattice,
    "lib/feq.R": install_feq,
    "lib/boundary.R": install_boundary,
    "lib/d3q19.R": install_d3q19,
    "lib/FMT.R": install_fmt,
    "lib/utils.R": install_utils,
}


def install(I):
    """bind the hooks; utils.R is what conf.R always loads, so its functions are available from the start"""
    for name, fn in HOOKS.items():
        I.source_hooks[name] = fn
    install_utils(I, I.globals)
