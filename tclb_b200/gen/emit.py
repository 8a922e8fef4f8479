"""IR -> source of one model library (CUDA sm_100a kernels + host restatement).

What the reference spreads over the templated translation unit ``src/cuda.cu.Rt`` (Node_Run,
:184-329), ``src/LatticeAccess.inc.cpp.Rt`` (pop/push/load, :260-507) and
``src/LatticeContainer.inc.cpp.Rt`` (executors and launches, :201-342, quantity kernels
:437-473, colour kernel :355-423) is emitted here as ONE fused kernel per (stage, globals
mode): pull-load -> node physics -> store (+ direct ghost stores into the neighbour
subdomain, possibly on a peer GPU) -> block reduction of Globals -> halo sequence flag.
The same node code is compiled for the host as the CPU oracle.
"""
from __future__ import annotations

import os
import re
from pathlib import Path
from typing import Dict, List

from .modeldef import ModelDef, nicename

_FLOAT_LIT = re.compile(
    r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)(?![\w.])")


def to_single_precision(code: str) -> str:
    """double literals -> float literals (the reference does the same with sed:
    src/makefile.main.Rt:293-296)."""
    out = []
    for line in code.split("\n"):
        if line.lstrip().startswith("#"):       # keep preprocessor lines (e.g. #define pi) as they are
            out.append(_FLOAT_LIT.sub(r"\1f", line) if line.lstrip().startswith("#define") else line)
        else:
            out.append(_FLOAT_LIT.sub(r"\1f", line))
    return "\n".join(out)


def _cstr(s) -> str:
    if s is None:
        return "0"
    return '"' + str(s).replace("\\", "\\\\").replace('"', '\\"').replace("\n", "\\n") + '"'


def _member_decls(names: List[str]) -> List[str]:
    """f[0]..f[8] -> real_t f[9]; scalars as they are (order of first appearance)."""
    arrays: Dict[str, int] = {}
    order: List[str] = []
    for n in names:
        m = re.match(r"^(\w+)\[(\d+)\]$", n)
        if m:
            base, i = m.group(1), int(m.group(2))
            if base not in arrays:
                order.append(base)
            arrays[base] = max(arrays.get(base, 0), i + 1)
        else:
            if n not in order:
                order.append(n)
    out = []
    for n in order:
        out.append(f"real_t {n}[{arrays[n]}];" if n in arrays else f"real_t {n};")
    return out


def _xidx(dx: int) -> str:
    """index expression of the node the density is pulled from (x wraps periodically)"""
    if dx == 0:
        return "i0"
    if dx == 1:
        return "im"
    if dx == -1:
        return "ip"
    return f"i0 - x + tclbrt::wrapx(x, {-dx}, a.nx)"


def _fmt_default(v) -> str:
    if isinstance(v, bool):
        return "1" if v else "0"
    if isinstance(v, (int, float)):
        return repr(float(v)) if isinstance(v, float) else str(v)
    return str(v)


def _lattice_tables(code: str) -> str:
    """Velocity/weight tables that model code may index at run time (the reference keeps them in constant memory,
    src/Global.h.Rt:173-213): emitted only when the code names them, as a __constant__ copy for the device pass and a
    plain array for the host pass of the same __host__ __device__ node code."""
    import numpy as np
    from .lbmlib import lattice_weights
    from .rlib import velocity_sets
    sets = velocity_sets()
    tabs = {}
    d2q9, d3q27 = sets["d2q9"], sets["d3q27"]
    tabs["wf"] = lattice_weights(d2q9[:, :2])
    tabs["wt"] = np.array([2 / 6] + [1 / 6] * 4)
    tabs["d2q9_ex"], tabs["d2q9_ey"] = d2q9[:, 0], d2q9[:, 1]
    tabs["d2q5_ex"], tabs["d2q5_ey"] = d2q9[:5, 0], d2q9[:5, 1]
    w27 = lattice_weights(d3q27)
    tabs["wg"] = w27
    tabs["wtConv"] = np.array(sorted(w27.tolist(), key=lambda w: -w))       # rest, axes, edges, corners
    tabs["wh"] = np.array([2 / 9] + [1 / 9] * 6 + [1 / 72] * 8)
    tabs["d3q27_ex"], tabs["d3q27_ey"], tabs["d3q27_ez"] = d3q27[:, 0], d3q27[:, 1], d3q27[:, 2]
    out = []
    for name, v in tabs.items():
        if re.search(r"(?<![\w.])" + name + r"\s*\[", code) is None:
            continue
        vals = ", ".join(repr(float(x)) for x in v)
        # base_t, not real_t: constants carry no tangents, and a __constant__ array of dual numbers cannot be initialised
        out.append(f"#ifdef __CUDACC__\nstatic __constant__ base_t {name}_dev_[{len(v)}] = {{{vals}}};\n#endif\n"
                   f"static const base_t {name}_host_[{len(v)}] = {{{vals}}};\n"
                   f"#ifdef __CUDA_ARCH__\n#define {name} {name}_dev_\n#else\n#define {name} {name}_host_\n#endif")
    return "\n".join(out)


# Particle access for model code (reference src/Particle.hpp:1-150 + the SolidAll container, src/SolidAll.h): a range
# over the particles whose padded bounding box contains the node; `for (auto p : SyncParticleIterator(X,Y,Z))` gets one
# object per particle with p.diff, p.cvel, p.rad, p.dist, p.applyForce(f); force and moment are added to the particle
# row when the object goes out of scope.  The container is the brute-force one (every particle is tested), which is what
# the reference configures by default; accumulation is one atomic per component and contributing node.
_PARTICLES = r"""
#define RFI_DATA_R 0
#define RFI_DATA_POS 1
#define RFI_DATA_VEL 4
#define RFI_DATA_ANGVEL 7
#define RFI_DATA_FORCE 10
#define RFI_DATA_MOMENT 13
#define RFI_DATA_SIZE 16
#define PART_MAR_VALUE @PART_MAR@
#define PART_MAR ((real_t)PART_MAR_VALUE)
TCLB_HD void particle_atomic_add(base_t *p, real_t v_) {
    const base_t v = tclb_value(v_);
#ifdef __CUDA_ARCH__
    atomicAdd(p, v);
#else
#pragma omp atomic
    *p += v;
#endif
}
struct Particle {
    vector_t cvel, diff;
    real_t rad, dist;
    TCLB_HD bool in() const { return dist < rad; }
};
struct ParticleI : Particle {
    base_t *row;
    TCLB_HD ParticleI(base_t *row_, real_t x, real_t y, real_t z) : row(row_) {
        rad = row[RFI_DATA_R];
        diff.x = x - row[RFI_DATA_POS + 0]; diff.y = y - row[RFI_DATA_POS + 1]; diff.z = z - row[RFI_DATA_POS + 2];
        dist = sqrt(diff.x * diff.x + diff.y * diff.y + diff.z * diff.z);
        const real_t ox = row[RFI_DATA_ANGVEL + 0], oy = row[RFI_DATA_ANGVEL + 1], oz = row[RFI_DATA_ANGVEL + 2];
        cvel.x = row[RFI_DATA_VEL + 0] + oy * diff.z - oz * diff.y;
        cvel.y = row[RFI_DATA_VEL + 1] + oz * diff.x - ox * diff.z;
        cvel.z = row[RFI_DATA_VEL + 2] + ox * diff.y - oy * diff.x;
    }
};
struct ParticleS : ParticleI {
    vector_t force, moment;
    bool touched;
    TCLB_HD ParticleS(base_t *row_, real_t x, real_t y, real_t z) : ParticleI(row_, x, y, z), touched(false) {
        force.x = force.y = force.z = 0;
        moment.x = moment.y = moment.z = 0;
    }
    TCLB_HD void applyForce(vector_t f) {
        force.x += f.x; force.y += f.y; force.z += f.z;
        moment.x -= f.y * diff.z - f.z * diff.y;
        moment.y -= f.z * diff.x - f.x * diff.z;
        moment.z -= f.x * diff.y - f.y * diff.x;
        touched = true;
    }
    TCLB_HD ~ParticleS() {
        if (!touched) return;
        particle_atomic_add(row + RFI_DATA_FORCE + 0, force.x); particle_atomic_add(row + RFI_DATA_FORCE + 1, force.y);
        particle_atomic_add(row + RFI_DATA_FORCE + 2, force.z);
        particle_atomic_add(row + RFI_DATA_MOMENT + 0, moment.x); particle_atomic_add(row + RFI_DATA_MOMENT + 1, moment.y);
        particle_atomic_add(row + RFI_DATA_MOMENT + 2, moment.z);
    }
};
// Range over the particles that can touch the node at (x,y,z): through the uniform-grid container when the runtime built
// one (reference SolidGrid: the two cells per axis around the node, `depth` slots each), else over every particle
// (reference SolidAll).  Either way a particle is handed out only if its padded bounding box contains the node.
TCLB_HD float tclb_word_float(int w) {
#ifdef __CUDA_ARCH__
    return __int_as_float(w);
#else
    float f; memcpy(&f, &w, 4); return f;
#endif
}
template <class P>
struct ParticleRange {
    base_t *data;
    int n;
    real_t x, y, z;
    const int *grid;
    int gn0, gn1, gn2, depth, lo0, lo1, lo2, hi0, hi1, hi2;
    TCLB_HD ParticleRange(base_t *data_, int n_, real_t x_, real_t y_, real_t z_, const int *grid_, const int *gn, int depth_,
                          const double *org, double inv_delta) : data(data_), n(n_), x(x_), y(y_), z(z_), grid(grid_) {
        gn0 = gn[0]; gn1 = gn[1]; gn2 = gn[2]; depth = depth_;
        lo0 = lo1 = lo2 = hi0 = hi1 = hi2 = 0;
        if (grid && depth >= 0) {
            const double p[3] = {(double)tclb_value(x), (double)tclb_value(y), (double)tclb_value(z)};
            int lo[3], hi[3];
            for (int k = 0; k < 3; k++) {
                const double c = (p[k] - org[k]) * inv_delta;
                lo[k] = (int)floor(c - 0.5); hi[k] = (int)floor(c + 0.5);
                if (lo[k] < 0) lo[k] = 0;
                if (hi[k] >= gn[k]) hi[k] = gn[k] - 1;
            }
            lo0 = lo[0]; lo1 = lo[1]; lo2 = lo[2]; hi0 = hi[0]; hi1 = hi[1]; hi2 = hi[2];
        }
    }
    TCLB_HD bool near(int i) const {
        const base_t *row = data + (size_t)i * RFI_DATA_SIZE;
        const real_t rr = row[RFI_DATA_R] + PART_MAR;
        return fabs(x - row[RFI_DATA_POS + 0]) <= rr && fabs(y - row[RFI_DATA_POS + 1]) <= rr && fabs(z - row[RFI_DATA_POS + 2]) <= rr;
    }
    struct It {
        const ParticleRange *r;
        int i;                 // particle index, r->n = end
        int c0, c1, c2, d;     // grid walk: cell and slot
        TCLB_HD void skip() {
            if (!r->grid) { while (i < r->n && !r->near(i)) i++; return; }
            if (r->depth < 0) {
                // threaded box tree (csrc/core/particle_tree.cpp): c0 = node to look at next; children follow their parent,
                // word 6 of a node is where to continue when its subtree is left out
                const float px = (float)tclb_value(r->x), py = (float)tclb_value(r->y), pz = (float)tclb_value(r->z);
                while (c0 < r->gn0) {
                    const int *nd = r->grid + (size_t)c0 * 8;
                    const bool in = px >= tclb_word_float(nd[0]) && py >= tclb_word_float(nd[1]) && pz >= tclb_word_float(nd[2]) &&
                                    px <= tclb_word_float(nd[3]) && py <= tclb_word_float(nd[4]) && pz <= tclb_word_float(nd[5]);
                    if (!in) { c0 = nd[6]; continue; }
                    const int k = nd[7];
                    if (k < 0) { c0++; continue; }
                    c0 = nd[6];
                    if (r->near(k)) { i = k; return; }
                }
                i = r->n;
                return;
            }
            for (; c2 <= r->hi2; c2++, c1 = r->lo1) {
                for (; c1 <= r->hi1; c1++, c0 = r->lo0) {
                    for (; c0 <= r->hi0; c0++, d = 0) {
                        const int *cell = r->grid + (((size_t)c2 * r->gn1 + c1) * r->gn0 + c0) * r->depth;
                        for (; d < r->depth; d++) {
                            const int k = cell[d];
                            if (k < 0) break;             // slots fill from the front
                            if (r->near(k)) { i = k; return; }
                        }
                    }
                }
            }
            i = r->n;
        }
        TCLB_HD P operator*() const { return P(r->data + (size_t)i * RFI_DATA_SIZE, r->x, r->y, r->z); }
        TCLB_HD It &operator++() { if (!r->grid) i++; else if (r->depth >= 0) d++; skip(); return *this; }
        TCLB_HD bool operator!=(const It &o) const { return i != o.i; }
    };
    TCLB_HD It begin() const {
        It it{this, 0, lo0, lo1, lo2, 0};
        if (grid && depth < 0) { it.c0 = 0; if (n == 0) it.i = n; else it.skip(); return it; }
        if (n == 0 || (grid && (lo0 > hi0 || lo1 > hi1 || lo2 > hi2))) it.i = n; else it.skip();
        return it;
    }
    TCLB_HD It end() const { return It{this, n, 0, 0, 0, 0}; }
};
// the older iterator form of model code: for (ParticleIterator p(X,Y,Z); p; ++p) { ... p.diff ... p.applyForce(f); }
// (rewritten by the generator to OldParticleIterator<...> p(*this, X,Y,Z)); the object is the range, its iterator and
// the current particle at once
struct ParticleDummyRow {
    base_t zero_row[RFI_DATA_SIZE];
    TCLB_HD ParticleDummyRow() { for (int i = 0; i < RFI_DATA_SIZE; i++) zero_row[i] = 0; }
};
template <class P>
struct ParticleRangeHolder {
    ParticleRange<P> range;
    typename ParticleRange<P>::It it;
    TCLB_HD ParticleRangeHolder(const ParticleRange<P> &r) : range(r), it(range.begin()) {}
    TCLB_HD base_t *row(base_t *dummy) const { return it != range.end() ? range.data + (size_t)it.i * RFI_DATA_SIZE : dummy; }
};
template <class P, class N>
struct OldParticleIterator : ParticleDummyRow, ParticleRangeHolder<P>, P {
    typedef ParticleRangeHolder<P> H;
    TCLB_HD OldParticleIterator(const N &node, real_t x, real_t y, real_t z)
        : ParticleDummyRow(), H(node.template particle_range<P>(x, y, z)), P(H::row(this->zero_row), x, y, z) {}
    TCLB_HD operator bool() const { return H::it != H::range.end(); }
    TCLB_HD OldParticleIterator &operator++() {
        static_cast<P *>(this)->~P();              // a ParticleS hands its force to the table here
        ++H::it;
        new (static_cast<P *>(this)) P(H::row(this->zero_row), H::range.x, H::range.y, H::range.z);
        return *this;
    }
};
"""


def _autosym_resolve(m: ModelDef, s, fidx) -> str:
    """what follows the plain pull of a stage under option autosym (reference SymmetryAccess::pop,
    src/LatticeAccess.inc.cpp.Rt:449-475 and resolve.symmetries :229-256)"""
    mode = int(getattr(m, "autosym", 0) or 0)
    if not mode:
        return ""
    L = []
    axes = "XYZ"
    have = {n.group for n in m.nodetypes}
    loaded = [d for d in m.densities if d.load.get(s.name) and not d.adjoint]
    if mode == 1:
        # mirror plane through the node: populations that would come from beyond it are the mirror partners held by this node
        for ax in range(3):
            if f"SYM{axes[ax]}" not in have:
                continue
            comp = lambda d: (d.dx, d.dy, d.dz)[ax]
            plus = [d for d in loaded if comp(d) < 0]
            minus = [d for d in loaded if comp(d) > 0]
            if not plus and not minus:
                continue
            L.append(f"    switch (n.n_type & NODE_SYM{axes[ax]}) {{")
            for tag, sel in (("plus", plus), ("minus", minus)):
                if sel:
                    L.append(f"    case NODE_{m.autosym_name}{axes[ax]}_{tag}:")
                    for d in sel:
                        p = d.sym[ax]
                        L.append(f"        n.{d.name} = {'-' if p.startswith('-') else ''}n.{p.lstrip('-')};")
                    L.append("        break;")
            L.append("    }")
    else:
        # mirror plane on the cell edge: on a symmetry node every population is pulled through the redirecting accessor
        groups = " | ".join(f"NODE_SYM{axes[ax]}" for ax in range(3) if f"SYM{axes[ax]}" in have)
        if groups:
            L.append(f"    if (n.n_type & ({groups})) {{")
            for d in loaded:
                L.append(f"        n.{d.name} = n.load_field({fidx[d.field]}, {-d.dx}, {-d.dy}, {-d.dz});")
            L.append("    }")
    return "\n".join(L)


# node code of a few older models reads the synthetic-turbulence time wave number straight from the reference's constant
# container (`constContainer.ST.TimeWN`, models/heat/experimental/d3q27_cumulant_heat); the value travels as the frequency of
# a trailing zero-amplitude mode (csrc/handlers/handlers.cpp, AcSyntheticTurbulence)
_CONST_CONTAINER = """    struct ConstContainerST_ { real_t TimeWN; };
    struct ConstContainer_ { ConstContainerST_ ST; };
    TCLB_HD ConstContainer_ const_container_() const {
        ConstContainer_ c;
        c.ST.TimeWN = ka_.st_n > 0 ? (real_t)ka_.st[ka_.st_n - 1].freq : (real_t)0;
        return c;
    }
#define constContainer const_container_()
    // the one-argument call of the same generation of model files: the field at this node
    TCLB_HD vector_t SyntheticTurbulence(real_t) const { return SyntheticTurbulence(X, Y, Z); }
"""


def _int_assignments_for_dual(code: str) -> str:
    """Adjoint source only.  Model code may keep a field expression in an INTEGER variable
    (`int solidFlag[27]; solidFlag[k] = PhaseF(1,0,0)/-998;`): an implicit real -> int conversion, which the dual-number type
    does not offer on purpose (its conversions are explicit, so that no arithmetic silently loses its tangents).  Where the
    model itself asks for an integer the truncation is the model's meaning (and its derivative is zero), so assignments to
    variables declared `int` get the conversion written out.  The ordinary source is left as it is."""
    names = set()
    for mt in re.finditer(r"\bint\s+([A-Za-z_]\w*)\s*(\[[^\]]*\])?\s*(?=[;=,)])", code):
        names.add(mt.group(1))
    names -= {"i", "j", "k"} & set()          # loop counters are handled like the rest: (int)tclb_value(0) is 0
    if not names:
        return code
    alt = "|".join(sorted(map(re.escape, names), key=len, reverse=True))
    pat = re.compile(r"(?<![\w.>])(" + alt + r")(\s*\[[^\]]*\])?\s*=(?!=)\s*([^;{}]+);")

    def fix(mt):
        rhs = mt.group(3).strip()
        if re.fullmatch(r"[-+]?\d+", rhs):     # plain integer literal: nothing to convert
            return mt.group(0)
        return f"{mt.group(1)}{mt.group(2) or ''} = (int)tclb_value({rhs});"
    return pat.sub(fix, code)


def _guard_type_switches(code: str, m: ModelDef) -> str:
    """`switch (NodeType & NODE_<GROUP>) { case NODE_<type>: ... }` statements of the model code get a test in front that,
    in the persistent kernels only (TCLB_NODE_PERSISTENT is a compile-time constant of the Node instance), skips the
    dispatch for nodes without a bit of that group: the arms' register shuffles then stay off the common path (measured:
    ring kernel of d3q27_cumulant +6 %, d3q19 +1 %; the same test in the one-block-per-tile kernels costs 20 %, hence the
    restriction).  Only switches whose every label is a node type of that group with a non-zero value and that have no
    `default:` are touched, so nothing changes in meaning.  The collision group is left alone (almost every node has it)."""
    if os.environ.get("TCLB_SWITCH_GUARD", "1") == "0":
        return code
    groups = {g.name: g for g in m.nodetypegroups}
    values = {n.name: (n.group, int(n.value)) for n in m.nodetypes}
    out, pos = [], 0
    for mt in re.finditer(r"switch\s*\(\s*NodeType\s*&\s*NODE_(\w+)\s*\)\s*\{", code):
        g = mt.group(1)
        if g not in groups or g in ("COLLISION", "SETTINGZONE") or mt.start() < pos:
            continue
        depth, i = 1, mt.end()
        while i < len(code) and depth > 0:
            depth += {"{": 1, "}": -1}.get(code[i], 0)
            i += 1
        body = code[mt.end():i - 1]
        # labels of this switch only (nested switches are rare in node-type dispatch; any doubt -> leave it)
        if re.search(r"\bdefault\s*:", body) or "switch" in body:
            continue
        labels = re.findall(r"\bcase\s+([^:]+?)\s*:", body)
        ok = bool(labels)
        for lab in labels:
            mm = re.fullmatch(r"NODE_(\w+)", lab.strip())
            if not mm or mm.group(1) not in values or values[mm.group(1)][0] != g or values[mm.group(1)][1] == 0:
                ok = False
        if not ok:
            continue
        out.append(code[pos:mt.start()])
        out.append(f"if (!TCLB_NODE_PERSISTENT || (NodeType & NODE_{g}) != 0) ")
        out.append(code[mt.start():i])
        pos = i
    out.append(code[pos:])
    return "".join(out)


def emit_model_source(m: ModelDef, libname: str, adjoint: bool = False) -> str:
    """Return the text of the generated .cu file (precision chosen by -DTCLB_DOUBLE).  With `adjoint` the source also
    carries the section compiled under -DTCLB_AD (the dual-number adjoint, see adjoint.cu.in); the ordinary source leaves
    it out so that work on the adjoint does not touch the primal libraries."""
    m.finalize()
    nf, nd = len(m.fields), len(m.densities)
    ns, nzs, ng, nq = len(m.settings), len(m.zonesettings), len(m.globals), len(m.quantities)
    gy = max([max(-f.miny, f.maxy) for f in m.fields] + [0])
    gz = max([max(-f.minz, f.maxz) for f in m.fields] + [0])
    gx = max([max(-f.minx, f.maxx) for f in m.fields] + [0])
    flag_t = "uint16_t" if m.flag_bits == 16 else "uint32_t"
    fidx = {f.name: i for i, f in enumerate(m.fields)}
    obj_idx = [i for i, g in enumerate(m.globals) if g.name == "Objective"][0]

    L: List[str] = []
    A = L.append
    A(f"// GENERATED by tclb_b200.gen.emit -- model {m.name} ({libname}); do not edit.")
    # fp32 divisions as reciprocal + multiply unless the model's tuning table (or TCLB_PREC_DIV=1) asks for the IEEE sequence
    # (measured, 512^3 fp32: reference cumulant files 20.2 -> 25.7 GLUPS, reference d3q19 35.5 -> 38.0, pf_velocity 8.2 -> 9.5)
    if not (getattr(m, "tuning", {}) or {}).get("prec_div", False) and os.environ.get("TCLB_PREC_DIV", "0") != "1":
        A("// nvcc-flags: -prec-div=false")
    A('#include <new>\n#include "tclb_node_rt.h"')
    A("#ifdef TCLB_DOUBLE\ntypedef double base_t;\n#define TCLB_REAL_NAME \"double\"\n#else\ntypedef float base_t;\n#define TCLB_REAL_NAME \"float\"\n#endif")
    # adjoint build (-DTCLB_AD, csrc/kernels/tclb_dual.h): the node code computes in dual numbers, memory stays base_t
    A("#ifdef TCLB_AD\n#include \"tclb_dual.h\"\ntypedef Dual<base_t, TCLB_AD_DIRS> real_t;\n#else\ntypedef base_t real_t;\n#endif")
    A("typedef tclb_vec3<real_t> vector_t;")
    # Color() of a model assigns real_t expressions to a float2; the adjoint build never calls it, so there the
    # members just swallow what they are given
    if adjoint:
        A("#ifdef TCLB_AD\nstruct tclb_ad_sink {\n"
          "    template <class U> TCLB_HD tclb_ad_sink &operator=(const U &) { return *this; }\n"
          "    template <class U> TCLB_HD tclb_ad_sink &operator+=(const U &) { return *this; }\n"
          "    template <class U> TCLB_HD tclb_ad_sink &operator-=(const U &) { return *this; }\n"
          "    template <class U> TCLB_HD tclb_ad_sink &operator*=(const U &) { return *this; }\n"
          "    template <class U> TCLB_HD tclb_ad_sink &operator/=(const U &) { return *this; }\n"
          "    TCLB_HD operator float() const { return 0.f; }\n};\n"
          "struct tclb_ad_float2 { tclb_ad_sink x, y; };\n#define float2 tclb_ad_float2\n#endif")
    # storage type of the lattice (reference ./configure --with-storage=half|half-shift|float|float-shift|double,
    # src/LatticeAccess.inc.cpp.Rt:14-35): what lives in HBM may be narrower than what the node code computes in, and a
    # per-field constant (AddDensity(shift=)) may be subtracted before storing so that small deviations keep their bits
    A(r"""#if defined(TCLB_STORAGE_HALF)
#include <cuda_fp16.h>
typedef __half storage_t;
#define TCLB_STORAGE_NAME "half"
TCLB_HD real_t storage_to_real(storage_t v) { return (real_t)__half2float(v); }
TCLB_HD storage_t real_to_storage(real_t v) { return __float2half((float)v); }
#elif defined(TCLB_STORAGE_FLOAT)
typedef float storage_t;
#define TCLB_STORAGE_NAME "float"
#define storage_to_real(v__) ((real_t)(v__))
#define real_to_storage(v__) ((float)(v__))
#else
typedef base_t storage_t;
#define TCLB_STORAGE_NAME TCLB_REAL_NAME
#define storage_to_real(v__) (v__)
#define real_to_storage(v__) (v__)
#endif
#ifdef TCLB_STORAGE_SHIFT
#define TCLB_LD(v__, s__) (storage_to_real(v__) + (real_t)(s__))
#define TCLB_ST(v__, s__) real_to_storage((v__) - (real_t)(s__))
#define TCLB_STORAGE_SHIFTED 1
#else
#define TCLB_LD(v__, s__) storage_to_real(v__)
#define TCLB_ST(v__, s__) real_to_storage(v__)
#define TCLB_STORAGE_SHIFTED 0
#endif""")
    A(f"typedef {flag_t} flag_t;")
    A("#ifdef _OPENMP\n#include <omp.h>\n#endif")
    tun = getattr(m, "tuning", {}) or {}
    A("#ifndef TCLB_MIN_BLOCKS\n#ifdef TCLB_DOUBLE\n#define TCLB_MIN_BLOCKS %d\n#else\n#define TCLB_MIN_BLOCKS %d\n#endif\n#endif"
      % (int(tun.get("min_blocks_double", 0)), int(tun.get("min_blocks_float", 0))))
    if tun.get("min_blocks_g_float") or tun.get("min_blocks_g_double"):
        # separate bound for the kernel variants that also reduce Globals (they need a few more registers)
        A("#ifndef TCLB_MIN_BLOCKS_G\n#ifdef TCLB_DOUBLE\n#define TCLB_MIN_BLOCKS_G %d\n#else\n#define TCLB_MIN_BLOCKS_G %d\n#endif\n#endif"
          % (int(tun.get("min_blocks_g_double", tun.get("min_blocks_double", 0))), int(tun.get("min_blocks_g_float", tun.get("min_blocks_float", 0)))))
    # which stage kernel a launch uses unless TCLB_KERNEL says otherwise (0 one block per tile, 1 persistent, 2 persistent +
    # bulk-copy ring), separately for launches without / with Globals, and the occupancy the persistent kernels are built for
    A("#ifndef TCLB_KERNEL_DEFAULT\n#define TCLB_KERNEL_DEFAULT %d\n#endif" % int(tun.get("kernel", 0)))
    A("#ifndef TCLB_KERNEL_DEFAULT_G\n#define TCLB_KERNEL_DEFAULT_G %d\n#endif" % int(tun.get("kernel_g", 2)))
    # interior fast path of the GPU stage kernels: a warp whose nodes carry no bits outside the collision and zone groups runs
    # a copy of the stage in which those bits are known to be zero (the boundary dispatch and its register shuffles fold away)
    plain_mask = 0
    for g in m.nodetypegroups:
        if g.name not in ("COLLISION", "SETTINGZONE"):
            plain_mask |= int(g.mask)
    A("#ifndef TCLB_RING_SLOTS\n#define TCLB_RING_SLOTS %d\n#endif" % int(tun.get("ring_slots", 1)))
    A("#ifndef TCLB_ELECT\n#define TCLB_ELECT %d\n#endif" % int(tun.get("elect", 0)))
    A("// true inside the node code of the persistent kernels (compile-time constant of the Node<GMODE> instance)")
    A("#define TCLB_NODE_PERSISTENT ((GMODE & 8) != 0)")
    A("#ifndef TCLB_FUSED_PUBLISH\n#define TCLB_FUSED_PUBLISH %d\n#endif" % int(tun.get("fused_publish", 1)))
    A("#define TCLB_PLAIN_MASK %du" % plain_mask)
    A("#ifndef TCLB_FASTPATH\n#define TCLB_FASTPATH %d\n#endif" % int(tun.get("fastpath", 0)))
    A("#ifndef TCLB_PERSIST_MIN_BLOCKS\n#ifdef TCLB_DOUBLE\n#define TCLB_PERSIST_MIN_BLOCKS %d\n#else\n#define TCLB_PERSIST_MIN_BLOCKS %d\n#endif\n#endif"
      % (int(tun.get("persist_min_blocks_double", 3)), int(tun.get("persist_min_blocks_float", 5))))
    A("namespace {")
    A(f"constexpr int N_FIELDS={nf}, N_DENSITIES={nd}, N_SETTINGS={ns}, N_ZONESETTINGS={nzs}, N_GLOBALS={ng}, N_QUANTITIES={nq};")
    A(f"constexpr int SUM_GLOBALS={m.sum_globals}, IN_OBJ_OFFSET={m.in_obj_offset}, GLOBALS_Objective={obj_idx};")
    A(f"constexpr int ZONE_SHIFT={m.zone_shift}, ZONE_MAX={m.zone_max};")
    A(f"constexpr int GX={gx}, GY={gy}, GZ={gz};")
    A("constexpr int NO_CUT = 65535, CUT_MAX = 1000; // cut distances are uint16 (reference src/types.h:25-29)")
    if "cut_t" in m.code or "CUT_LEN" in m.code:
        # the type and the length macro of the reference's cut distances (src/types.h:25,29); a cut is the fraction of the link
        # times CUT_MAX here as there (the scale itself is this build's own: host side csrc/host/geometry.cpp calc_cut)
        A("typedef uint16_t cut_t;")
        A("#define CUT_LEN(x__) ((1.0f / CUT_MAX) * (x__))")
    for k, v in m.options.items():
        if v:
            A(f"#define OPTIONS_{k} {int(v) if not isinstance(v, bool) else 1}")
    for n in m.nodetypes:
        A(f"#define NODE_{n.name} {n.value}")
    for g in m.nodetypegroups:
        A(f"#define NODE_{g.name} {g.mask}")
    for i, f in enumerate(m.fields):
        A(f"#define FIELD_{nicename(f.name)} {i}")
    for i, s in enumerate(m.settings):
        A(f"#define SETTINGS_{s.name} {i}")
    for i, s in enumerate(m.zonesettings):
        A(f"#define ZONESETTINGS_{s.name} {i}")
    for i, g in enumerate(m.globals):
        A(f"#define GLOBALS_{g.name}_IDX {i}")

    A(r"""
struct KArgs {
    int nx, ny, nz, NY, NZ;
    long long fstride;
    int off_x, off_y, off_z, tot_nx, tot_ny, tot_nz;
    real_t pos_x, pos_y, pos_z;
    const storage_t *in;
    storage_t *out;
    // per-density source pointers, already shifted by the pull offset in y/z (x is handled by the wrapped
    // index), and per-field destination pointers: an access is ONE 32-bit-index IMAD.WIDE + LDG/STG
    const storage_t *ldp[N_DENSITIES];
    storage_t *stp[N_FIELDS];
    int obj_enabled;
    const flag_t *flags;
    const uint16_t *cuts;
    double *globals;
    double *gl_ring[TCLB_MAX_RANKS];
    unsigned *gl_counter;
    base_t *gl_partial;
    long long gl_partial_stride;
    unsigned gl_epoch;
    int gl_publish, gl_rank, gl_nranks, n_blocks;
    const base_t *zone;
    int zone_len, zone_index, zone_stride;
    int iter, reset_iter;
    storage_t *halo[9];
    int halo_ny[9], halo_nz[9];
    unsigned *sig_count;
    unsigned *sig_remote[9];
    const unsigned *sig_wait[9];
    unsigned sig_epoch;
    int sig_enable, n_border_blocks;
    int nxb, nyb, by, ybr;
    int rx, ry, rz, rnx, rny, rnz;
    void *qout;
    real_t qscale;
    const tclb_st_mode_t *st;
    int st_n;
    base_t *particles;           // [n_particles][16]: r, pos, vel, angular velocity, force, moment (RFI layout)
    int n_particles;
    const int *pgrid;            // uniform-grid container over the particles (NULL: scan all), see tclb_launch_t
    int pg_n[3], pg_depth;
    double pg_origin[3], pg_inv_delta;
    // launches captured in a CUDA graph take iteration and halo epoch from device memory (dyn[0], dyn[1] = values at
    // the start of the graph) plus a constant offset, so one instantiated graph can be replayed; NULL otherwise
    const int *dyn;
    int dyn_iter_off;
    unsigned dyn_epoch_off;
    // adjoint build, gradient with respect to zonal settings: tangent direction j is seeded where the node code reads
    // zonal setting ad_zs_set[j] in zone ad_zs_zone[j] (ad_zs_dt[j] = 0: the value, 1: its _DT time derivative); -1 = unused
    short ad_zs_set[16], ad_zs_zone[16], ad_zs_dt[16];
    real_t settings[N_SETTINGS > 0 ? N_SETTINGS : 1];
};
// Kernel variant bit 4 (GMODE & 4) = "launched from a captured graph": only that variant reads the device-resident
// counters.  Keeping the run-time test out of the ordinary variants matters: with it the d3q27_cumulant kernel
// lost 4 % (same instruction count, different schedule) - measured on B200.
template <int GMODE> TCLB_HD int k_iter(const KArgs &a) { return (GMODE & 4) ? a.dyn[0] + a.dyn_iter_off : a.iter; }
template <int GMODE> TCLB_HD unsigned k_epoch(const KArgs &a) { return (GMODE & 4) ? (unsigned)a.dyn[1] + a.dyn_epoch_off : a.sig_epoch; }
template <int GMODE> TCLB_HD int k_zone_index(const KArgs &a) {
    return ((GMODE & 4) && a.zone_len > 1) ? (a.dyn[0] + a.dyn_iter_off) % a.zone_len : a.zone_index;
}
""")
    # ---- macros visible to model code ----
    macros: List[tuple] = []
    def M(name, body):
        macros.append((name, body))
    # cell-centred real coordinates, like the reference's getX() (src/LatticeAccess.inc.cpp.Rt:274, src/Lattice.cu.Rt:66-71)
    M("X", "((real_t)(n_x + ka_.off_x) + ka_.pos_x)"); M("Y", "((real_t)(n_y + ka_.off_y) + ka_.pos_y)")
    M("Z", "((real_t)(n_z + ka_.off_z) + ka_.pos_z)")
    M("NodeType", "n_type")
    M("Time", "(k_iter<GMODE>(ka_))")
    M("average_iter", "(k_iter<GMODE>(ka_) - ka_.reset_iter)")
    # plain settings are const-reference members of the node (not macros: names like `z`, `T`, `dt` must not
    # rewrite `u.z` or a parameter called z); after inlining they fold into constant-bank operands
    # zonal settings are read where they are used (a load per node for values only boundary nodes need would
    # cost more than it saves); a setting the model code assigns to keeps a per-node member like the reference's
    def _assigned(name: str) -> bool:
        return re.search(r"(?<![\w.])" + re.escape(name) + r"\s*(=(?!=)|\+=|-=|\*=|/=)", m.code) is not None
    member_zs = {s.name for s in m.zonesettings if s.preload and _assigned(s.name)}
    for i, s in enumerate(m.zonesettings):
        if s.name not in member_zs:
            M(s.name, f"zone_setting({i})")
        M(f"{s.name}_DT", f"zone_setting_dt({i})")
    for i, g in enumerate(m.globals):
        fn = "AddToGlobal" if g.op == "SUM" else "MaxToGlobal"
        M(f"AddTo{g.name}(x__)", f"{fn}<{i}>(x__)")
    for n in m.nodetypes:
        gmask = [g.mask for g in m.nodetypegroups if g.name == n.group][0]
        M(f"Iam{n.name}", f"((n_type & {gmask}) == {n.value})")
    for g in m.nodetypegroups:
        M(f"Iam{g.name}", f"(n_type & {g.mask})")
    for i, f in enumerate(m.fields):
        M(f"{nicename(f.name)}(...)", f"load_field({i}, __VA_ARGS__)")
        M(f"{nicename(f.name)}_dyn(...)", f"load_field({i}, __VA_ARGS__)")
    qn = 0
    for c in range(3):
        for b in range(3):
            for a_ in range(3):
                if (a_, b, c) != (0, 0, 0):
                    M(f"Q{a_}{b}{c}", f"(getQ({qn}))")
                    qn += 1
    for name, body in macros:
        A(f"#define {name} {body}")
    A(_lattice_tables(m.code))
    A(_PARTICLES.replace("@PART_MAR@", repr(float(getattr(m, "part_margin", None) or 0.5))))
    # storage shift of every field (0 unless the model asked for one); folds to a constant after inlining
    A("TCLB_HD real_t field_shift(int fi) {\n    switch (fi) {")
    for i, f in enumerate(m.fields):
        if f.shift:
            A(f"    case {i}: return (real_t){float(f.shift)!r};")
    A("    default: return (real_t)0;\n    }\n}")

    # adjoint build: ids of the (field, offset) pairs a stage can read through stencil accessors
    st_fields = [(i, f) for i, f in enumerate(m.fields) if not f.simple_access]
    A("#ifdef TCLB_AD")
    A("TCLB_HD int ad_stencil_id(int fi, int dx, int dy, int dz) {\n    switch (fi) {")
    off = 0
    st_table = []
    for i, f in st_fields:
        ex, ey, ez = f.maxx - f.minx + 1, f.maxy - f.miny + 1, f.maxz - f.minz + 1
        A(f"    case {i}: return {off} + ((dz - ({f.minz})) * {ey} + (dy - ({f.miny}))) * {ex} + (dx - ({f.minx}));")
        for dz_ in range(f.minz, f.maxz + 1):
            for dy_ in range(f.miny, f.maxy + 1):
                for dx_ in range(f.minx, f.maxx + 1):
                    st_table.append((i, dx_, dy_, dz_))
        off += ex * ey * ez
    A("    default: return -(1 << 29);\n    }\n}")
    A(f"constexpr int AD_N_STENCIL = {len(st_table)};")
    A("#endif")
    autosym = int(getattr(m, "autosym", 0) or 0)
    A(f"#define TCLB_AUTOSYM {autosym}")
    if autosym:
        A(f"#define TCLB_AUTOSYM_SHIFT {int(m.autosym_shift)}")
        axes = "XYZ"
        def nt(name):
            r = [n for n in m.nodetypes if n.name == name]
            return r[0] if r else None
        for what, suf in (("mask", None), ("plus", "_plus"), ("minus", "_minus")):
            A(f"TCLB_HD unsigned t_sym_{what}(int ax) {{\n    switch (ax) {{")
            for ax in range(3):
                n_ = nt(f"{m.autosym_name}{axes[ax]}_plus")
                if n_ is None:
                    continue
                val = n_.mask if suf is None else nt(f"{m.autosym_name}{axes[ax]}{suf}").value
                A(f"    case {ax}: return {val}u;")
            A("    default: return 0u;\n    }\n}")
        A("TCLB_HD int t_sym_field(int fi, int ax) {\n    switch (fi * 3 + ax) {")
        for i, f in enumerate(m.fields):
            for ax in range(3):
                A(f"    case {i * 3 + ax}: return {fidx[f.sym[ax].lstrip('-')]};")
        A("    default: return fi;\n    }\n}")
        A("TCLB_HD real_t t_sym_sign(int fi, int ax) {\n    switch (fi * 3 + ax) {")
        for i, f in enumerate(m.fields):
            for ax in range(3):
                if f.sym[ax].startswith("-"):
                    A(f"    case {i * 3 + ax}: return (real_t)-1;")
        A("    default: return (real_t)1;\n    }\n}")
    members = _member_decls([d.name for d in m.densities] + [f.name for f in m.fields])
    A(r"""
#ifdef __CUDA_ARCH__
constexpr int GSTRIDE = 33;     // shared-memory strip [global][lane], padded against bank conflicts
#else
constexpr int GSTRIDE = 1;
#endif
template <int GMODE>
struct Node {
    const KArgs &ka_;
    int n_x, n_y, n_z;
    int n_idx;
    flag_t n_type;
    real_t *gacc_;        // per-node Globals accumulators: a strip of shared memory on the GPU (no registers held
                          // across the collision), a local array on the host; element i lives at gacc_[i * GSTRIDE]
#ifdef TCLB_AD
    int ad_lo_ = -(1 << 30);   // first stencil-read input id seeded in this pass (adjoint build), see load_field_plain
#endif
""")
    for d in members:
        A("    " + d)
    for s in m.zonesettings:
        if s.name in member_zs:
            A(f"    real_t {s.name};")
    for s in m.settings:
        A(f"    const real_t &{s.name};")
    # brace initialisers: a field of the same name is a function-like macro and would swallow `name(`
    init_refs = "".join(f", {s.name}{{a_.settings[{i}]}}" for i, s in enumerate(m.settings))
    # GMODE bit 8: the accumulators live across the nodes a thread visits (persistent kernel) and are cleared by the
    # kernel, not per node; `type` >= 0 hands in a node type that was already fetched (shared-memory ring)
    A("    TCLB_HD Node(const KArgs &a_, int x, int y, int z, real_t *gacc = nullptr, int type = -1) : ka_(a_), n_x(x), n_y(y), n_z(z), gacc_(gacc)" + init_refs + " {")
    A(r"""
        n_idx = ((z + GZ) * ka_.NY + (y + GY)) * ka_.nx + x;
        n_type = type >= 0 ? (flag_t)type : ka_.flags[n_idx];
        if (!(GMODE & 8)) {
            if (GMODE & 1) {
                for (int i = 0; i < N_GLOBALS; i++) gacc_[i * GSTRIDE] = (i >= SUM_GLOBALS) ? (real_t)-1e30 : (real_t)0;
            } else if ((GMODE & 3) == 2) gacc_[0] = 0;
        }""")
    for i, s in enumerate(m.zonesettings):
        if s.name in member_zs:
            A(f"        {s.name} = zone_setting({i});")
    A(r"""    }
#ifdef TCLB_AD
    TCLB_HD void ad_seed_zone(real_t &v, int s, int dt) const {
        const int zn = (int)(n_type >> ZONE_SHIFT);
        for (int j = 0; j < TCLB_AD_DIRS && j < 16; j++)
            if (ka_.ad_zs_set[j] == s && ka_.ad_zs_zone[j] == zn && ka_.ad_zs_dt[j] == dt) v.d[j] = 1;
    }
#endif
    TCLB_HD real_t zone_setting(int s) const {
        // the table is tiny (time levels x settings x zones): 32-bit index arithmetic
        real_t v = ka_.zone[(k_zone_index<GMODE>(ka_) * N_ZONESETTINGS + s) * ka_.zone_stride + (int)(n_type >> ZONE_SHIFT)];
#ifdef TCLB_AD
        ad_seed_zone(v, s, 0);
#endif
        return v;
    }
    TCLB_HD real_t zone_setting_dt(int s) const {
        const int zi = k_zone_index<GMODE>(ka_);
        int ip = zi + 1 < ka_.zone_len ? zi + 1 : 0;
        int im = zi > 0 ? zi - 1 : ka_.zone_len - 1;
        int zn = n_type >> ZONE_SHIFT;
        real_t v = (ka_.zone[((long long)ip * N_ZONESETTINGS + s) * ka_.zone_stride + zn] -
                    ka_.zone[((long long)im * N_ZONESETTINGS + s) * ka_.zone_stride + zn]) * (real_t)0.5;
#ifdef TCLB_AD
        ad_seed_zone(v, s, 1);
#endif
        return v;
    }
    TCLB_HD real_t load_field_plain(int fi, int dx, int dy, int dz) const {
        int xx = tclbrt::wrapx(n_x, dx, ka_.nx);
        long long i = ((long long)(n_z + dz + GZ) * ka_.NY + (n_y + dy + GY)) * ka_.nx + xx;
        real_t v = TCLB_LD(ka_.in[fi * ka_.fstride + i], field_shift(fi));
#ifdef TCLB_AD
        // adjoint build: a read through a stencil accessor is an input of the stage like a pulled density; the directions
        // of the current chunk that stand for (field, offset) pairs are ad_lo_ .. ad_lo_ + TCLB_AD_DIRS - 1
        const int j = ad_stencil_id(fi, dx, dy, dz) - ad_lo_;
        if (j >= 0 && j < TCLB_AD_DIRS) v.d[j] = 1;
#endif
        return v;
    }
#if TCLB_AUTOSYM > 0
    // option autosym (reference SymmetryAccess, src/LatticeAccess.inc.cpp.Rt:376-447): a read that crosses a mirror plane
    // of this node (node types Symmetry<X|Y|Z>_<plus|minus>) is redirected to the mirror partner of the field at the
    // reflected offset; z is resolved first, then y, then x, as in the reference's load_ -> loadY_ -> loadX_ chain
    TCLB_HD real_t load_field(int fi, int dx, int dy, int dz) const {
        int d[3] = {dx, dy, dz};
        real_t sg = 1;
#pragma unroll
        for (int ax = 2; ax >= 0; ax--) {
            const unsigned grp = (unsigned)n_type & t_sym_mask(ax);
            if (d[ax] > 0 && grp != 0 && grp == t_sym_plus(ax)) { d[ax] = TCLB_AUTOSYM_SHIFT - d[ax]; sg *= t_sym_sign(fi, ax); fi = t_sym_field(fi, ax); }
            else if (d[ax] < 0 && grp != 0 && grp == t_sym_minus(ax)) { d[ax] = -TCLB_AUTOSYM_SHIFT - d[ax]; sg *= t_sym_sign(fi, ax); fi = t_sym_field(fi, ax); }
        }
        return sg * load_field_plain(fi, d[0], d[1], d[2]);
    }
#else
    TCLB_HD real_t load_field(int fi, int dx, int dy, int dz) const { return load_field_plain(fi, dx, dy, dz); }
#endif
    TCLB_HD real_t load_field(int fi, int dx, int dy) const { return load_field(fi, dx, dy, 0); }
    // offsets computed in the arithmetic type (e.g. a wall normal): truncated like the implicit conversion of the primal build
    template <class A_, class B_, class C_>
    TCLB_HD real_t load_field(int fi, A_ dx, B_ dy, C_ dz) const { return load_field(fi, (int)tclb_value(dx), (int)tclb_value(dy), (int)tclb_value(dz)); }
    template <class P> TCLB_HD ParticleRange<P> particle_range(real_t x, real_t y, real_t z) const {
        return ParticleRange<P>(ka_.particles, ka_.n_particles, x, y, z, ka_.pgrid, ka_.pg_n, ka_.pg_depth, ka_.pg_origin, ka_.pg_inv_delta);
    }
    TCLB_HD ParticleRange<ParticleS> SyncParticleIterator(real_t x, real_t y, real_t z) const { return particle_range<ParticleS>(x, y, z); }
    TCLB_HD ParticleRange<ParticleI> ParticleIterator(real_t x, real_t y, real_t z) const { return particle_range<ParticleI>(x, y, z); }
    TCLB_HD uint16_t getQ(int d) const {
        return ka_.cuts ? ka_.cuts[(long long)d * ((long long)ka_.nx * ka_.NY * ka_.NZ) + n_idx] : (uint16_t)NO_CUT;
    }
@CONST_CONTAINER@    TCLB_HD vector_t SyntheticTurbulence(real_t x, real_t y, real_t z) const {
        vector_t v; v.x = 0; v.y = 0; v.z = 0;
        for (int i = 0; i < ka_.st_n; i++) {
            const tclb_st_mode_t &w = ka_.st[i];
            real_t ph = (real_t)(w.k[0] * x + w.k[1] * y + w.k[2] * z + w.freq * (double)k_iter<GMODE>(ka_));
            real_t s = sin(ph), c = cos(ph);
            v.x += (real_t)w.a[0] * c + (real_t)w.b[0] * s;
            v.y += (real_t)w.a[1] * c + (real_t)w.b[1] * s;
            v.z += (real_t)w.a[2] * c + (real_t)w.b[2] * s;
        }
        return v;
    }
    // Globals: SUM globals also add (zone's <Global>InObj weight) * x to Objective.  The reference's line
    // (src/cuda.cu.Rt:113-119) adds the weight WITHOUT the factor x, which makes Objective independent of the
    // flow; the product is what its optimisation handlers document and what the adjoint here differentiates.
    // GMODE bits: 1 = accumulate every Global, 2 = add the InObj weights to Objective.  The launcher picks the
    // variant without bit 2 when every weight is zero (the usual case), so the hot kernel carries no weight loads.
    template <int I> TCLB_HD void AddToGlobal(real_t x) {
        if (GMODE & 1) gacc_[I * GSTRIDE] += x;
        if ((GMODE & 2) && I < SUM_GLOBALS && I != GLOBALS_Objective)
            gacc_[((GMODE & 1) ? GLOBALS_Objective : 0) * GSTRIDE] += zone_setting(I + IN_OBJ_OFFSET) * x;
    }
    template <int I> TCLB_HD void MaxToGlobal(real_t x) {
        if (GMODE & 1) gacc_[I * GSTRIDE] = tclbrt::tmax(gacc_[I * GSTRIDE], x);
    }
// ---------------------------------------------------------------- model physics ----
#ifdef TCLB_DOUBLE
""".replace("@CONST_CONTAINER@", _CONST_CONTAINER if "constContainer" in m.code else ""))
    # older iterator form of some shipped models: `for (ParticleIterator p(X,Y,Z); p; ++p)` declares the iterator as an object
    # (also the Tree.../Full... spellings of earlier container generations)
    code = re.sub(r"\bSync(?:Full)?ParticleIterator\s+(\w+)\s*\(", r"OldParticleIterator<ParticleS, Node<GMODE> > \1(*this, ", m.code)
    code = re.sub(r"(?<![<\w])(?:Tree|Full)?ParticleIterator\s+(\w+)\s*\(", r"OldParticleIterator<ParticleI, Node<GMODE> > \1(*this, ", code)
    code = _guard_type_switches(code, m)
    if adjoint:
        code = _int_assignments_for_dual(code)
    A(code)
    A("#else")
    A(to_single_precision(code))
    A("#endif")
    A("// ---------------------------------------------------------------- end of model physics ----")
    # quantity accessors dispatch
    A("    template <int Q> TCLB_HD void quantity(real_t *o) {")
    for i, q in enumerate(m.quantities):
        if q.vector:
            A(f"        if (Q == {i}) {{ vector_t v = get{q.name}(); o[0] = v.x; o[1] = v.y; o[2] = v.z; }}")
        else:
            A(f"        if (Q == {i}) {{ o[0] = get{q.name}(); }}")
    A("    }")
    A("    template <int S> TCLB_HD void run_stage() {")
    for i, s in enumerate(m.stages):
        A(f"        if (S == {i}) {s.main}();")
    A("    }")
    A("};")
    for name, _ in macros:
        A(f"#undef {name.split('(')[0]}")

    # ---- load / store per stage ----
    A(r"""
TCLB_HD long long src_index(const KArgs &a, int x, int y, int z, int dx, int dy, int dz) {
    int xx = x - dx;
    if (dx != 0) { if (xx < 0) xx += a.nx; if (xx >= a.nx) xx -= a.nx; }
    return ((long long)(z - dz + GZ) * a.NY + (y - dy + GY)) * a.nx + xx;
}

// store one value into every neighbour ghost that will read it (reach known at compile time)
template <int MINY, int MAXY, int MINZ, int MAXZ>
TCLB_HD void halo_put(const KArgs &a, int fi, int x, int y, int z, storage_t v) {
#pragma unroll
    for (int sz = -1; sz <= 1; sz++) {
#pragma unroll
        for (int sy = -1; sy <= 1; sy++) {
            if (sy == 0 && sz == 0) continue;
            if (sy > 0 && !(MINY < 0 && y >= a.ny + MINY)) continue;
            if (sy < 0 && !(MAXY > 0 && y < MAXY)) continue;
            if (sz > 0 && !(MINZ < 0 && z >= a.nz + MINZ)) continue;
            if (sz < 0 && !(MAXZ > 0 && z < MAXZ)) continue;
            const int d = (sy + 1) + 3 * (sz + 1);
            storage_t *dst = a.halo[d];
            if (!dst) continue;
            const int yy = sy > 0 ? y - a.ny : (sy < 0 ? y + a.halo_ny[d] : y);
            const int zz = sz > 0 ? z - a.nz : (sz < 0 ? z + a.halo_nz[d] : z);
            const long long NYd = a.halo_ny[d] + 2 * GY, NZd = a.halo_nz[d] + 2 * GZ;
            dst[(long long)fi * (NYd * NZd * a.nx) + ((long long)(zz + GZ) * NYd + (yy + GY)) * a.nx + x] = v;
        }
    }
}
""")
    shift_of = {f.name: float(f.shift or 0.0) for f in m.fields}
    for si, s in enumerate(m.stages):
        A(f"// stage {si}: {s.name} (main {s.main})")
        A(f"template <int GMODE> TCLB_HD void load_stage_{si}(Node<GMODE> &n, const KArgs &a, int x, int i0, int im, int ip) {{")
        for di, d in enumerate(m.densities):
            if d.load.get(s.name):
                A(f"    n.{d.name} = TCLB_LD(a.ldp[{di}][{_xidx(d.dx)}], {shift_of[d.field]!r});")
            elif d.default is not None:
                A(f"    n.{d.name} = (real_t){d.default!r};")
        A(_autosym_resolve(m, s, fidx))
        A("}")
        A(f"template <int GMODE> TCLB_HD void store_stage_{si}(Node<GMODE> &n, const KArgs &a, int x, int y, int z, bool border) {{")
        for f in m.fields:
            if f.save.get(s.name):
                fi = fidx[f.name]
                A(f"    a.stp[{fi}][n.n_idx] = TCLB_ST(n.{f.name}, {shift_of[f.name]!r});")
        A("    if (border) {")
        for f in m.fields:
            if f.save.get(s.name) and (f.miny or f.maxy or f.minz or f.maxz):
                fi = fidx[f.name]
                A(f"        halo_put<{f.miny},{f.maxy},{f.minz},{f.maxz}>(a, {fi}, x, y, z, TCLB_ST(n.{f.name}, {shift_of[f.name]!r}));")
        A("    }")
        A("}")
    # ---- persistent kernel with a shared-memory ring fed by bulk asynchronous copies (scaffold.cu.in k_stage_persist) ----
    # Per stage: every loaded density is one row of the tile (TCLB_TILE_X consecutive x at the pulled y/z), copied by ONE
    # cp.async.bulk; the |dx| elements a pulled row needs from beyond the tile (periodic in x) come as single-element
    # cp.async copies, so a row sits in shared memory with 16 bytes of room on either side.
    A("#if defined(__CUDACC__) && !defined(TCLB_AD)")
    A("constexpr int TCLB_TILE_X = 128;")
    A("constexpr int RING_PAD = 16 / (int)sizeof(storage_t) > 0 ? 16 / (int)sizeof(storage_t) : 1;   // elements of room before/after a row")
    A("constexpr int RING_ROW = (TCLB_TILE_X + 2 * RING_PAD) * (int)sizeof(storage_t);               // bytes per row")
    ring_ok = []
    for si, s_ in enumerate(m.stages):
        ld = [(di, d) for di, d in enumerate(m.densities) if d.load.get(s_.name)]
        edges = []      # (row, element position relative to the row start, density index)
        for k, (di, d) in enumerate(ld):
            if d.dx > 0:
                edges += [(k, -j, di) for j in range(1, d.dx + 1)]
            elif d.dx < 0:
                edges += [(k, 128 + j, di) for j in range(0, -d.dx)]
        ok = bool(ld) and all(abs(d.dx) <= 2 for _, d in ld) and not int(getattr(m, "autosym", 0) or 0)
        ring_ok.append(ok)
        A(f"constexpr int RING_ROWS_{si} = {len(ld) if ok else 0}, RING_EDGES_{si} = {len(edges) if ok else 0};")
        A(f"constexpr int RING_BYTES_{si} = RING_ROWS_{si} * RING_ROW + TCLB_TILE_X * (int)sizeof(flag_t);")
        A(f"constexpr bool RING_OK_{si} = {'true' if ok else 'false'} && sizeof(storage_t) >= 4;")
        # chosen by default (launches with Globals) only where it was measured to win (512^3 fp32, Globals every iteration,
        # one block per tile -> ring): d3q27_cumulant 24.1 -> 26.9 GLUPS, d3q27_cumulant_AVG (34 densities) 13.4 -> 16.9, d3q19
        # 32.5 -> 37.4; it loses on small velocity sets (d2q9_bc 72.5 -> 55.9, d2q9_kuper 20.6 -> 17.0: too little arithmetic
        # per tile to hide the ring's bookkeeping) and on d3q27_pf_velocity's 50-density stages (8.2 -> 5.5)
        A(f"constexpr bool RING_DEFAULT_{si} = RING_OK_{si} && {'true' if (ok and 15 <= len(ld) <= 40) else 'false'};")
        A(f"template <int GMODE> TCLB_D void load_stage_ring_{si}(Node<GMODE> &n, const unsigned char *sb, int t) {{")
        if ok:
            for k, (di, d) in enumerate(ld):
                A(f"    n.{d.name} = TCLB_LD(((const storage_t *)(sb + {k} * RING_ROW))[RING_PAD + t - ({d.dx})], {shift_of[d.field]!r});")
            for di, d in enumerate(m.densities):
                if not d.load.get(s_.name) and d.default is not None:
                    A(f"    n.{d.name} = (real_t){d.default!r};")
        A("}")
        # rows: lane 0 of every warp issues its share (row k belongs to warp k % 4), so no warp is held up by the whole
        # list; warp 0 also announces the byte count.  `row0` = index of the tile's first node in a field
        A(f"TCLB_D void ring_issue_rows_{si}(const KArgs &a, unsigned char *sb, void *bar, long long row0, int wid) {{")
        if ok:
            # one switch on the warp number (not one test per row: the tests were 6 % of the kernel's instructions)
            rows = [(k, f"a.ldp[{di}] + row0", "TCLB_TILE_X * (unsigned)sizeof(storage_t)", f"{k} * RING_ROW + 16") for k, (di, d) in enumerate(ld)]
            rows.append((len(ld), "a.flags + row0", "TCLB_TILE_X * (unsigned)sizeof(flag_t)", f"RING_ROWS_{si} * RING_ROW"))
            A("    switch (wid) {")
            for w in range(4):
                A(f"    case {w}:")
                if w == 0:
                    A(f"        tclbrt::mbar_arrive_expect_tx(bar, (unsigned)(RING_ROWS_{si} * TCLB_TILE_X * (int)sizeof(storage_t) + TCLB_TILE_X * (int)sizeof(flag_t)));")
                for k, src, nbytes, dst in rows:
                    if k % 4 == w:
                        A(f"        tclbrt::bulk_g2s(sb + {dst}, {src}, {nbytes}, bar);")
                A("        break;")
            A("    }")
        A("}")
        if ok and edges:
            # one packed word per edge element in constant memory: row | field << 8 | (position + 8) << 16 | (dy + 2) << 24 |
            # (dz + 2) << 27 (the source pointer is rebuilt from these: indexing the kernel parameter array with a lane
            # dependent index would make the compiler copy it to local memory)
            def _w(e):
                d_ = m.densities[e[2]]
                return e[0] | (fidx[d_.field] << 8) | ((e[1] + 8) << 16) | ((d_.dy + 2) << 24) | ((d_.dz + 2) << 27)
            A(f"__constant__ int t_ring_edges_{si}[{len(edges)}] = {{{', '.join(str(_w(e)) for e in edges)}}};")
        A(f"TCLB_D void ring_issue_edges_{si}(const KArgs &a, unsigned char *sb, int lane, long long row0, int x0) {{")
        if ok and edges:
            A(f"    for (int e = lane; e < {len(edges)}; e += 32) {{")
            A(f"        const int w = t_ring_edges_{si}[e], row = w & 255, fi = (w >> 8) & 255, p = ((w >> 16) & 255) - 8, dy = ((w >> 24) & 7) - 2, dz = ((w >> 27) & 7) - 2;")
            A("        int xs = x0 + p; if (xs < 0) xs += a.nx; if (xs >= a.nx) xs -= a.nx;")
            A("        const storage_t *src = a.in + (long long)fi * a.fstride - ((long long)dz * a.NY + dy) * a.nx + (row0 - x0) + xs;")
            A("        tclbrt::cp_async_elem<(int)sizeof(storage_t)>(sb + row * RING_ROW + 16 + p * (int)sizeof(storage_t), src);")
            A("    }")
        A("}")
    A("template <int S, int GMODE> TCLB_D void load_stage_ring(Node<GMODE> &n, const unsigned char *sb, int t) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) load_stage_ring_{si}(n, sb, t);")
    A("}")
    A("template <int S> TCLB_D void ring_issue_rows(const KArgs &a, unsigned char *sb, void *bar, long long row0, int wid) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) ring_issue_rows_{si}(a, sb, bar, row0, wid);")
    A("}")
    A("template <int S> TCLB_D void ring_issue_edges(const KArgs &a, unsigned char *sb, int lane, long long row0, int x0) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) ring_issue_edges_{si}(a, sb, lane, row0, x0);")
    A("}")
    A("template <int S> constexpr bool ring_ok() { return " + " || ".join(f"(S == {si} && RING_OK_{si} && RING_BYTES_{si} * TCLB_RING_SLOTS <= 100 * 1024)" for si in range(len(m.stages))) + "; }")
    A("template <int S> constexpr bool ring_default() { return ring_ok<S>() && (" + " || ".join(f"(S == {si} && RING_DEFAULT_{si})" for si in range(len(m.stages))) + "); }")
    A("template <int S> constexpr int ring_bytes() { return " + " + ".join(f"(S == {si} ? RING_BYTES_{si} : 0)" for si in range(len(m.stages))) + "; }")
    A("#endif")
    A("template <int GMODE> TCLB_HD void load_all(Node<GMODE> &n, const KArgs &a, int x, int i0, int im, int ip) {")
    for di, d in enumerate(m.densities):
        A(f"    n.{d.name} = TCLB_LD(a.ldp[{di}][{_xidx(d.dx)}], {shift_of[d.field]!r});")
    class _AllStage:       # quantities see every density, resolved at symmetry nodes like a stage that loads them all
        name = "\0all"
    for d in m.densities:
        d.load["\0all"] = True
    A(_autosym_resolve(m, _AllStage, fidx))
    for d in m.densities:
        d.load.pop("\0all", None)
    A("}")
    A("template <int S, int GMODE> TCLB_HD void load_stage(Node<GMODE> &n, const KArgs &a, int x, int i0, int im, int ip) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) load_stage_{si}(n, a, x, i0, im, ip);")
    A("}")
    A("const signed char t_dens_dy[] = {" + ", ".join(str(d.dy) for d in m.densities) + "};")
    A("const signed char t_dens_dz[] = {" + ", ".join(str(d.dz) for d in m.densities) + "};")
    A("const int t_dens_field[] = {" + ", ".join(str(fidx[d.field]) for d in m.densities) + "};")
    A("template <int S, int GMODE> TCLB_HD void store_stage(Node<GMODE> &n, const KArgs &a, int x, int y, int z, bool border) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) store_stage_{si}(n, a, x, y, z, border);")
    A("}")

    A(_SCAFFOLD)

    # ---- adjoint build: seeds, collection and scatter per stage (see adjoint.cu.in) ----
    A("#ifdef TCLB_AD")
    ad_mark = len(L)
    A("TCLB_HD void ad_add(base_t *p, base_t v);")
    A("TCLB_HD void ad_put(base_t *p, base_t v) { if (v != (base_t)0) *p = v; }")
    loaded = [[(di, d) for di, d in enumerate(m.densities) if d.load.get(s.name)] for s in m.stages]
    A("TCLB_HD int ad_inputs_of_stage(int s) {\n    switch (s) {")
    for si in range(len(m.stages)):
        A(f"    case {si}: return {len(loaded[si])} + AD_N_STENCIL;")
    A("    default: return 0;\n    }\n}")
    A("TCLB_HD int ad_loaded_of_stage(int s) {\n    switch (s) {")
    for si in range(len(m.stages)):
        A(f"    case {si}: return {len(loaded[si])};")
    A("    default: return 0;\n    }\n}")
    # scatter of the stencil inputs: id -> (field, offset)
    A("TCLB_HD void ad_scatter_stencil(const KArgs &a, base_t *adj_in, int x, int y, int z, int id, base_t g) {")
    A("    if (g == (base_t)0) return;")
    A("    int fi = 0, dx = 0, dy = 0, dz = 0;\n    switch (id) {")
    for k, (i, dx_, dy_, dz_) in enumerate(st_table):
        A(f"    case {k}: fi = {i}; dx = {dx_}; dy = {dy_}; dz = {dz_}; break;")
    A("    default: return;\n    }")
    A("    const int xx = tclbrt::wrapx(x, dx, a.nx);")
    A("    ad_add(adj_in + (long long)fi * a.fstride + ((long long)(z + dz + GZ) * a.NY + (y + dy + GY)) * a.nx + xx, g);")
    A("}")
    for si, s in enumerate(m.stages):
        A(f"template <int GMODE> TCLB_HD void ad_seed_stage_{si}(Node<GMODE> &n, int c) {{")
        for k, (di, d) in enumerate(loaded[si]):
            A(f"    if (c == {k} / TCLB_AD_DIRS) n.{d.name}.d[{k} % TCLB_AD_DIRS] = 1;")
        A("}")
        A(f"template <int GMODE> TCLB_HD real_t ad_collect_stage_{si}(Node<GMODE> &n, const KArgs &a, const base_t *adj_out) {{")
        A("    real_t g = 0;")
        for f in m.fields:
            if f.save.get(s.name):
                A(f"    g += n.{f.name} * adj_out[{fidx[f.name]} * a.fstride + n.n_idx];")
        A("    return g;\n}")
        A(f"TCLB_HD void ad_scatter_stage_{si}(const KArgs &a, base_t *adj_in, int x, int i0, int im, int ip, int c, const real_t &g) {{")
        readers = {}
        for _, d in loaded[si]:
            readers[d.field] = readers.get(d.field, 0) + 1
        for k, (di, d) in enumerate(loaded[si]):
            # a field pulled by ONE density has one reader per cell (the node at cell + c): a plain store into the
            # zero-filled buffer; several densities on one field (stencil reads) need the atomic
            put = "ad_put" if readers[d.field] == 1 else "ad_add"
            A(f"    if (c == {k} / TCLB_AD_DIRS) {put}(adj_in + (a.ldp[{di}] - a.in) + ({_xidx(d.dx)}), g.d[{k} % TCLB_AD_DIRS]);")
        A("}")
    A("template <int S, int GMODE> TCLB_HD void ad_seed_stage(Node<GMODE> &n, int c) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) ad_seed_stage_{si}(n, c);")
    A("}")
    A("template <int S, int GMODE> TCLB_HD real_t ad_collect_stage(Node<GMODE> &n, const KArgs &a, const base_t *adj_out) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) return ad_collect_stage_{si}(n, a, adj_out);")
    A("    return real_t(0);\n}")
    A("template <int S> TCLB_HD void ad_scatter_stage(const KArgs &a, base_t *adj_in, int x, int i0, int im, int ip, int c, const real_t &g) {")
    for si in range(len(m.stages)):
        A(f"    if (S == {si}) ad_scatter_stage_{si}(a, adj_in, x, i0, im, ip, c, g);")
    A("}")
    A(_AD_SCAFFOLD)
    A("}  // namespace")
    A('extern "C" int tclb_model_adjoint_stage(int stage, const tclb_launch_t *L, const void *adj_out, void *adj_in) {')
    A("    KArgs a; int bx, by; fill_args(a, L, bx, by);")
    A("    switch (stage) {")
    for si in range(len(m.stages)):
        A(f"    case {si}: return run_adjoint_stage<{si}>(a, L, (const base_t *)adj_out, (base_t *)adj_in);")
    A('    default: set_error("bad stage"); return -1;\n    }\n}')
    A('extern "C" int tclb_model_adjoint_settings(int stage, const tclb_launch_t *L, const void *adj_out, const int *which, int n, double *grad) {')
    A("    KArgs a; int bx, by; fill_args(a, L, bx, by);")
    A("    switch (stage) {")
    for si in range(len(m.stages)):
        A(f"    case {si}: return run_adjoint_settings<{si}>(a, L, (const base_t *)adj_out, which, n, grad);")
    A('    default: set_error("bad stage"); return -1;\n    }\n}')
    A('extern "C" int tclb_model_adjoint_zsettings(int stage, const tclb_launch_t *L, const void *adj_out, const int *zs, const int *zone, const int *dt, int n, double *grad) {')
    A("    KArgs a; int bx, by; fill_args(a, L, bx, by);")
    A("    switch (stage) {")
    for si in range(len(m.stages)):
        A(f"    case {si}: return run_adjoint_zsettings<{si}>(a, L, (const base_t *)adj_out, zs, zone, dt, n, grad);")
    A('    default: set_error("bad stage"); return -1;\n    }\n}')
    A('extern "C" const char *tclb_model_adjoint_error(void) { return g_error; }')
    A(f'extern "C" int tclb_model_adjoint_info(int *real_size, int *dirs, int *n_fields) {{ *real_size = (int)sizeof(base_t); *dirs = TCLB_AD_DIRS; *n_fields = N_FIELDS; return TCLB_ABI_VERSION; }}')
    if not adjoint:
        del L[ad_mark:]
        A('#error "this source was generated without the adjoint section (emit_model_source(..., adjoint=True))"')
    A("#else  // !TCLB_AD")

    # ---- dispatch tables ----
    A("int run_stage_impl(int stage, const tclb_launch_t *L) {")
    A("    KArgs a; int bx, by; fill_args(a, L, bx, by);")
    A("    // kernel variant: 0 none, 1 all Globals, 2 Objective weights only, 3 all Globals + Objective weights")
    A("    // 4 = no Globals, iteration/epoch from device memory (launches captured in a CUDA graph)")
    A("    int gm = L->globals_mode == TCLB_GLOBALS_INTEGRATE ? (L->obj_enabled ? 3 : 1) : L->globals_mode;")
    A('    if (L->dyn) { if (gm != 0) { set_error("graph launches carry no Globals"); return -1; } gm = 4; }')
    A("    switch (stage * 8 + gm) {")
    for si in range(len(m.stages)):
        for gm in range(5):
            A(f"    case {si * 8 + gm}: return launch_stage<{si}, {gm}>(a, L, bx, by);")
    A('    default: set_error("bad stage/globals mode"); return -1;')
    A("    }\n}")
    A("int get_quantity_impl(int q, const tclb_launch_t *L) {")
    A("    KArgs a; int bx, by; fill_args(a, L, bx, by);")
    A("    switch (q) {")
    for i, q in enumerate(m.quantities):
        A(f"    case {i}: return launch_quantity<{i}, {3 if q.vector else 1}>(a, L);")
    A('    default: set_error("bad quantity"); return -1;')
    A("    }\n}")
    A("int kernel_info_impl(int stage, int gmode, int *regs, int *smem, int *lmem, int *maxthr) {")
    A("#ifdef __CUDACC__")
    A("    cudaFuncAttributes at; cudaError_t e = cudaErrorInvalidValue;")
    A("    switch (stage * 4 + gmode) {")
    for si in range(len(m.stages)):
        for gm in range(4):
            A(f"    case {si * 4 + gm}: e = cudaFuncGetAttributes(&at, k_stage<{si}, {gm}>); break;")
    A("    }")
    A("    if (e != cudaSuccess) { cudaGetLastError(); return -1; }")
    A("    *regs = at.numRegs; *smem = (int)at.sharedSizeBytes; *lmem = (int)at.localSizeBytes; *maxthr = at.maxThreadsPerBlock; return 0;")
    A("#else\n    return -1;\n#endif\n}")

    # ---- metadata ----
    for s in m.settings + m.zonesettings:
        if s.derived is not None:
            A(f"double derive_{s.name}(double {s.name}) {{ return {s.equation}; }}")
    A("const tclb_density_t t_densities[] = {")
    for d in m.densities:
        A(f"    {{{_cstr(d.name)}, {_cstr(d.field)}, {_cstr(d.group)}, {_cstr(d.comment)}, {d.dx}, {d.dy}, {d.dz}, {fidx[d.field]}, "
          f"{int(d.parameter)}, {int(d.average)}, {int(d.adjoint)}, {int(d.default is not None)}, {float(d.default or 0.0)!r}}},")
    A("};")
    A("const tclb_field_t t_fields[] = {")
    for f in m.fields:
        A(f"    {{{_cstr(f.name)}, {_cstr(f.group)}, {_cstr(f.comment)}, {f.minx}, {f.maxx}, {f.miny}, {f.maxy}, {f.minz}, {f.maxz}, "
          f"{int(f.parameter)}, {int(f.average)}, {int(f.adjoint)}}},")
    A("};")
    A("const double t_field_shift[] = {" + ", ".join(repr(float(f.shift or 0.0)) for f in m.fields) + "};")
    sidx = {s.name: i for i, s in enumerate(m.settings)}
    def setting_rows(tab, zonal):
        for s in tab:
            tgt = sidx[s.derived] if s.derived is not None else -1
            fn = f"derive_{s.name}" if s.derived is not None else "0"
            A(f"    {{{_cstr(s.name)}, {_cstr(s.unit)}, {_cstr(_fmt_default(s.default))}, {_cstr(s.comment)}, {int(zonal)}, {tgt}, {fn}, {int(s.adjoint)}}},")
    A("const tclb_setting_t t_settings[] = {")
    setting_rows(m.settings, False)
    A("    {0,0,0,0,0,-1,0,0}};")
    A("const tclb_setting_t t_zonesettings[] = {")
    setting_rows(m.zonesettings, True)
    A("    {0,0,0,0,0,-1,0,0}};")
    A("const tclb_global_t t_globals[] = {")
    for g in m.globals:
        A(f"    {{{_cstr(g.name)}, {_cstr(g.unit)}, {_cstr(g.comment)}, {0 if g.op == 'SUM' else 1}, {int(g.adjoint)}, {g.base!r}}},")
    A("};")
    A("const tclb_quantity_t t_quantities[] = {")
    for q in m.quantities:
        A(f"    {{{_cstr(q.name)}, {_cstr(q.unit)}, {_cstr(q.comment)}, {int(q.vector)}, {int(q.adjoint)}}},")
    A("    {0,0,0,0,0}};")
    A("const tclb_nodetype_t t_nodetypes[] = {")
    for n in m.nodetypes:
        A(f"    {{{_cstr(n.name)}, {_cstr(n.group)}, {n.value}u, {n.mask}u, {n.shift}, {n.bits}, {int(n.save)}}},")
    A("};")
    A("const tclb_nodegroup_t t_nodegroups[] = {")
    for g in m.nodetypegroups:
        # capacity is an int in the ABI: the catch-all group of a 32-bit flag word holds 2^32 - 1, reported as INT_MAX
        A(f"    {{{_cstr(g.name)}, {g.mask}u, {g.shift}, {g.bits}, {min(int(g.capacity), 2147483647)}, {int(g.save)}}},")
    A("};")
    stage_index = {s.name: i for i, s in enumerate(m.stages)}
    for si, s in enumerate(m.stages):
        A(f"const unsigned char t_stage{si}_load[] = {{{', '.join(str(int(bool(d.load.get(s.name)))) for d in m.densities)}}};")
        A(f"const unsigned char t_stage{si}_save[] = {{{', '.join(str(int(bool(f.save.get(s.name)))) for f in m.fields)}}};")
        A(f"const unsigned char t_stage{si}_read[] = {{{', '.join(str(int(bool(f.read.get(s.name)))) for f in m.fields)}}};")
    A("const tclb_stage_t t_stages[] = {")
    for si, s in enumerate(m.stages):
        A(f"    {{{_cstr(s.name)}, {_cstr(s.main)}, {int(s.particle)}, t_stage{si}_load, t_stage{si}_save, t_stage{si}_read}},")
    A("};")
    for ai, a_ in enumerate(m.actions):
        A(f"const int t_action{ai}_stages[] = {{{', '.join(str(stage_index[s]) for s in a_.stages)}}};")
    A("const tclb_action_t t_actions[] = {")
    for ai, a_ in enumerate(m.actions):
        A(f"    {{{_cstr(a_.name)}, {len(a_.stages)}, t_action{ai}_stages}},")
    A("};")
    A("const tclb_option_t t_options[] = {")
    for k, v in m.options.items():
        A(f"    {{{_cstr(k)}, {int(v)}}},")
    A("    {0, 0}};")
    # objectives: every Global by name, plus the model's AddObjective expressions of the Globals
    from .poly import Poly as _Poly, to_c as _to_c
    objs = {g.name: _Poly.sym(g.name) for g in m.globals}
    for k, v in (getattr(m, "objectives", {}) or {}).items():
        objs[k] = _Poly.lift(v[0] if hasattr(v, "__len__") and not isinstance(v, (str, _Poly)) else v)
    gnames = [g.name for g in m.globals]
    for k, e in objs.items():
        used = sorted({s_ for mono in e.t for s_, _ in mono})
        bad = [u for u in used if u not in gnames]
        if bad:
            raise ValueError(f"objective {k} uses {bad}, which are not Globals of the model")
        A(f"static void OF_{k}(const double *glob, double *obj, double *grad) {{")
        for u in used:
            A(f"    const double {u} = glob[{gnames.index(u)}];")
        A(f"    obj[0] = {_to_c(e)};")
        for gi, gname in enumerate(gnames):
            A(f"    grad[{gi}] = {_to_c(e.deriv(gname))};")
        A("}")
    A("const tclb_objective_t t_objectives[] = {")
    for k in objs:
        A(f"    {{{_cstr(k)}, OF_{k}}},")
    A("    {0, 0}};")
    A(f"""
const tclb_model_t t_model = {{
    TCLB_ABI_VERSION, {_cstr(libname)}, {_cstr(m.name)}, (int)sizeof(real_t), (int)sizeof(flag_t),
    ZONE_SHIFT, ZONE_MAX, SUM_GLOBALS, IN_OBJ_OFFSET,
    N_DENSITIES, t_densities, N_FIELDS, t_fields, N_SETTINGS, t_settings, N_ZONESETTINGS, t_zonesettings,
    N_GLOBALS, t_globals, N_QUANTITIES, t_quantities,
    {len(m.nodetypes)}, t_nodetypes, {len(m.nodetypegroups)}, t_nodegroups,
    {len(m.stages)}, t_stages, {len(m.actions)}, t_actions, {len(m.options)}, t_options,
    GY, GZ,
    run_stage_impl, get_quantity_impl, color_impl, kernel_info_impl, last_error_impl,
    (int)sizeof(storage_t), TCLB_STORAGE_SHIFTED, TCLB_STORAGE_NAME, t_field_shift,
    (double)PART_MAR_VALUE,
    {len(objs)}, t_objectives
}};
}}  // namespace
extern "C" const tclb_model_t *tclb_get_model(void) {{ return &t_model; }}
extern "C" long long tclb_model_launch_count(void) {{ return g_launch_count; }}
#endif  // !TCLB_AD
""")
    return "\n".join(L) + "\n"


_SCAFFOLD = (Path(__file__).with_name("scaffold.cu.in")).read_text()
_AD_SCAFFOLD = (Path(__file__).with_name("adjoint.cu.in")).read_text()
