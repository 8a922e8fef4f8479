"""Single-lattice flow models built from the generator library:

* ``d2q9_SRT``  - BGK tutorial model (reference models/flow/d2q9_SRT: Dynamics.R + Dynamics.c, 215 lines)
* ``d3q27``     - d3q27 MRT with Smagorinsky LES and drag/flux Globals (reference models/flow/d3q27)
* ``d2q9_les`` / ``d3q19_les`` - Smagorinsky variants of the MRT models (reference models/flow/d2q9_les, d3q19_les)

The declared surface (densities, settings, globals, node types, quantities) follows the reference
models; the physics is generated from :mod:`tclb_b200.gen.lbmlib`.
"""
from __future__ import annotations

import numpy as np

from ..gen.lbmlib import (bounce_index, d2q9_velocities, d3q19_velocities, full_bounce_back, lattice_weights, mrt_eq,
                          symmetry, sym_moments, sym_populations, tensor_velocities, zouhe)
from ..gen.modeldef import ModelDef
from ..gen.poly import PV, Poly, PVec, c_assign, psum, to_c
from . import ModelSpec


def _moments(f, U):
    d = U.shape[1]
    return [f @ U[:, k].astype(float) for k in range(d)]


# ------------------------------------------------------------------------------------------ d2q9_SRT
def build_d2q9_srt(options=None) -> ModelDef:
    m = ModelDef("d2q9_SRT", dict(options or {}))
    U = d2q9_velocities()
    for i, u in enumerate(U):
        m.AddDensity(name=f"f[{i}]", dx=int(u[0]), dy=int(u[1]))
    m.AddQuantity(name="U", unit="m/s", vector=True)
    m.AddQuantity(name="Rho", unit="kg/m3")
    m.AddSetting(name="omega", comment="inverse of relaxation time")
    m.AddSetting(name="nu", omega="1.0/(3*nu+0.5)", default=0.16666666, comment="viscosity")
    m.AddSetting(name="Velocity", default=0, comment="inlet/outlet/init velocity", zonal=True)
    m.AddSetting(name="Velocity_x", default=0, comment="inlet/outlet/init velocity in x", zonal=True)
    m.AddSetting(name="Velocity_y", default=0, comment="inlet/outlet/init velocity in y", zonal=True)
    m.AddSetting(name="GravitationX", default=0, comment="body/external acceleration", zonal=True)
    m.AddSetting(name="GravitationY", default=0, comment="body/external acceleration", zonal=True)
    m.AddSetting(name="Density", default=1, comment="Density")
    for n in ["EPressure", "WPressure", "WVelocity", "EVelocity", "Solid", "Wall"]:
        m.AddNodeType(name=n, group="BOUNDARY")
    m.AddNodeType(name="BGK", group="COLLISION")
    f = PV("f[", range(9), "]")
    # equilibrium in (d, u): feq(rho, J) with J = d*u
    EQ = mrt_eq(U, "d", ("Jx", "Jy"), orthogonal=False)
    feq = EQ.feq.subst({"Jx": Poly.sym("d") * Poly.sym("u.x"), "Jy": Poly.sym("d") * Poly.sym("u.y")})
    jx, jy = _moments(f, U)
    dens = m.density_dicts()
    c = []
    A = c.append
    A("""CudaDeviceFunction float2 Color() {
	float2 ret;
	vector_t u = getU();
	ret.x = sqrt(u.x*u.x + u.y*u.y);
	ret.y = 1;
	return ret;
}""")
    A("CudaDeviceFunction void SetEquilibrium(real_t f_eq[9], real_t d, vector_t u) {")
    A(c_assign(PV("f_eq[", range(9), "]"), feq))
    A("}")
    A("""CudaDeviceFunction void Init() {
	vector_t u;
	u.x = Velocity_x; u.y = Velocity_y; u.z = 0;
	real_t f_eq[9];
	SetEquilibrium(f_eq, Density, u);
	for (int i = 0; i < 9; i++) f[i] = f_eq[i];
}
CudaDeviceFunction void Run() {
	switch (NodeType & NODE_BOUNDARY) {
	case NODE_Solid:
	case NODE_Wall: BounceBack(); break;
	case NODE_EVelocity: EVelocity(); break;
	case NODE_WPressure: WPressure(); break;
	case NODE_WVelocity: WVelocity(); break;
	case NODE_EPressure: EPressure(); break;
	}
	if (NodeType & NODE_BGK) CollisionBGK();
}""")
    A(f"CudaDeviceFunction real_t getRho() {{\n\treturn {to_c(f.sum())};\n}}")
    A("CudaDeviceFunction vector_t getU() {")
    A(f"\treal_t d = {to_c(f.sum())};\n\tvector_t u;")
    A(f"\tu.x = ({to_c(jx)})/d + GravitationX*0.5;\n\tu.y = ({to_c(jy)})/d + GravitationY*0.5;\n\tu.z = 0;\n\treturn u;\n}}")
    A("CudaDeviceFunction void CollisionBGK() {\n\treal_t d = getRho();\n\tvector_t u;")
    # the body force enters as a velocity shift of the equilibrium (G/omega), like the tutorial model
    A(f"\tu.x = ({to_c(jx)})/d + GravitationX/omega;\n\tu.y = ({to_c(jy)})/d + GravitationY/omega;\n\tu.z = 0;")
    A("\treal_t f_eq[9];\n\tSetEquilibrium(f_eq, d, u);\n\tfor (int i = 0; i < 9; i++) f[i] = f[i] + omega*(f_eq[i]-f[i]);\n}")
    same_group = [dict(d, group="f") for d in dens]      # every density of this model is declared in a group of its own
    A("CudaDeviceFunction void BounceBack() {\n" + full_bounce_back(same_group, "f") + "}")
    # classic Zou-He faces in x (on-site, rho or u prescribed)
    EQj = mrt_eq(U, "rho", ("Jx", "Jy"), orthogonal=False)
    A("CudaDeviceFunction void EVelocity() {\n" + zouhe(EQj, f, 1, -1, "velocity", V="Velocity_x") + "}")
    A("CudaDeviceFunction void WVelocity() {\n" + zouhe(EQj, f, 1, 1, "velocity", V="Velocity_x") + "}")
    A("CudaDeviceFunction void WPressure() {\n\treal_t Jx, Jy, Jz, rho;\n\trho = Density;\n" +
      "\n".join(zouhe(EQj, f, 1, 1, "pressure", predefined=True).split("\n")[1:]) + "}")
    A("CudaDeviceFunction void EPressure() {\n\treal_t Jx, Jy, Jz, rho;\n\trho = Density;\n" +
      "\n".join(zouhe(EQj, f, 1, -1, "pressure", predefined=True).split("\n")[1:]) + "}")
    m.code = "\n".join(c)
    return m


# ------------------------------------------------------------------------------------------ generic MRT by moment order
def _orthogonal_mrt_collision(A, U, f, names_J, force_syms, smag=None, les_flag=None, extra=""):
    """relax the non-conserved moments of an integer-orthogonalised polynomial basis: second order with `omega`
    (optionally Smagorinsky-corrected), higher orders to equilibrium; force added to the momentum in between."""
    d = U.shape[1]
    EQ = mrt_eq(U, "rho", names_J, orthogonal=True)
    Q = len(f)
    R = PV("R", range(Q))
    Rn = R.copy()
    cons = np.nonzero(EQ.order < 2)[0]
    for i in cons:
        Rn[i] = EQ.Req[i]
    sel = np.nonzero(EQ.order > 1)[0]
    S = PVec([Poly.sym("om_r") if o == 2 else Poly.const(0) for o in EQ.order])
    A("\treal_t " + ", ".join(to_c(x) for x in Rn) + ";")
    A(sym_moments(f, EQ.mat, U, Rn))
    A(c_assign(Rn[sel], (Rn - EQ.Req)[sel]))
    A("\treal_t om_r = 1.0 - omega;")
    if smag:
        # |Pi_neq| from the second-order non-equilibrium moments (projected back), Hou et al. closure
        second = np.nonzero(EQ.order == 2)[0]
        norm = (EQ.mat ** 2).sum(axis=0)
        q = psum(Rn[i] * Rn[i] * float(1.0 / norm[i]) for i in second)
        A(f"\tif ({les_flag}) {{")
        A(f"\t\treal_t qn = sqrt({to_c(q)})*18.0*{smag}*{smag}/rho;")
        A("\t\treal_t tau0 = 1.0/omega;")
        A("\t\treal_t taut = 0.5*(sqrt(tau0*tau0 + qn) + tau0);")
        A("\t\tom_r = 1.0 - 1.0/taut;")
        A("\t}")
    A(c_assign(Rn[sel], (Rn * S)[sel]))
    A(extra)
    for k in range(d):
        A(f"\t{names_J[k]} = {names_J[k]} + rho*{force_syms[k]};")
    A(c_assign(Rn[sel], (Rn + EQ.Req)[sel]))
    A(sym_populations(Rn, np.linalg.inv(EQ.mat), U, f))
    return EQ


def build_d3q27(options=None) -> ModelDef:
    m = ModelDef("d3q27", dict(options or {}))
    # the reference orders the velocities as expand.grid(-1:1,-1:1,-1:1)
    U = np.array([[x, y, z] for z in (-1, 0, 1) for y in (-1, 0, 1) for x in (-1, 0, 1)], dtype=int)
    names = [f"f{i}" for i in range(27)]
    for i, (n, u) in enumerate(zip(names, U)):
        m.AddDensity(name=n, dx=int(u[0]), dy=int(u[1]), dz=int(u[2]), comment=f"density F {i}", group="f")
    m.AddQuantity(name="P", unit="Pa")
    m.AddQuantity(name="U", unit="m/s", vector=True)
    m.AddQuantity(name="Fd", unit="N", vector=True)
    m.AddSetting(name="omega", comment="One over relaxation time")
    m.AddSetting(name="nu", omega="1.0/(3*nu + 0.5)", default=0.16666666, comment="Viscosity")
    m.AddSetting(name="Velocity", default="0m/s", comment="Inlet velocity", zonal=True)
    m.AddSetting(name="Pressure", default="0Pa", comment="Inlet pressure", zonal=True)
    m.AddSetting(name="Smag", comment="Smagorinsky constant")
    m.AddSetting(name="Turbulence", comment="Turbulence intensity", zonal=True)
    for a in "XYZ":
        m.AddSetting(name=f"Force{a}", comment=f"Force force {a}")
    for a in "XYZ":
        m.AddGlobal(name=f"{a}Flux", comment="Volume flux", unit="m3/s")
    for a in "XYZ":
        m.AddGlobal(name=f"{a}DragForce", comment="Solid drag force", unit="N")
    m.AddNodeType(name="Smagorinsky", group="LES")
    m.AddNodeType(name="Stab", group="ENTROPIC")
    for n in ["NSymmetry", "SSymmetry", "ISymmetry", "EPressure", "Solid", "Wall", "WPressure", "WVelocity"]:
        m.AddNodeType(name=n, group="BOUNDARY")
    m.AddNodeType(name="MRT", group="COLLISION")
    f = PV(names)
    jx, jy, jz = _moments(f, U)
    dens = m.density_dicts("f")
    EQ = mrt_eq(U, "rho", ("Jx", "Jy", "Jz"), orthogonal=False)
    c = []
    A = c.append
    A(f"CudaDeviceFunction real_t getRho() {{\n\treturn {to_c(f.sum())};\n}}")
    A(f"CudaDeviceFunction real_t getP() {{\n\treturn (({to_c(f.sum())})-1.0)/3.0;\n}}")
    A("CudaDeviceFunction vector_t getU() {\n\treal_t d = getRho();\n\tvector_t u;")
    A(f"\tu.x = {to_c(jx)};\n\tu.y = {to_c(jy)};\n\tu.z = {to_c(jz)};")
    A("\tu.x = (u.x+ForceX*0.5)/d;\n\tu.y = (u.y+ForceY*0.5)/d;\n\tu.z = (u.z+ForceZ*0.5)/d;\n\treturn u;\n}")
    A("CudaDeviceFunction vector_t getFd() {\n\tvector_t force;\n\tif ((NodeType & NODE_BOUNDARY) == NODE_Wall) {")
    A(f"\t\tforce.x = {to_c(jx * -2.0)};\n\t\tforce.y = {to_c(jy * -2.0)};\n\t\tforce.z = {to_c(jz * -2.0)};")
    A("\t} else { force.x = 0; force.y = 0; force.z = 0; }\n\treturn force;\n}")
    A("""CudaDeviceFunction float2 Color() {
	float2 ret;
	vector_t u = getU();
	ret.x = sqrt(u.x*u.x + u.y*u.y + u.z*u.z);
	ret.y = (NodeType == NODE_Solid) ? 0 : 1;
	return ret;
}""")
    A("CudaDeviceFunction void BounceBack() {")
    A(full_bounce_back(dens, "f"))
    A("\tvector_t force;")
    A(f"\tforce.x = {to_c(jx * 2.0)};\n\tforce.y = {to_c(jy * 2.0)};\n\tforce.z = {to_c(jz * 2.0)};")
    A("\tAddToXDragForce(force.x);\n\tAddToYDragForce(force.y);\n\tAddToZDragForce(force.z);\n}")
    A("CudaDeviceFunction void WVelocity() {\n" + zouhe(EQ, f, 1, 1, "velocity") + "}")
    A("CudaDeviceFunction void WPressure() {\n" + zouhe(EQ, f, 1, 1, "pressure") + "}")
    A("CudaDeviceFunction void EPressure() {\n" + zouhe(EQ, f, 1, -1, "pressure") + "}")
    A("CudaDeviceFunction void NSymmetry() {\n" + symmetry(dens, 2, -1) + "}")
    A("CudaDeviceFunction void SSymmetry() {\n" + symmetry(dens, 2, 1) + "}")
    A("CudaDeviceFunction void ISymmetry() {\n" + symmetry(dens, 3, 1) + "}")
    A("""CudaDeviceFunction void Run() {
	switch (NodeType & NODE_BOUNDARY) {
	case NODE_WPressure: WPressure(); break;
	case NODE_WVelocity: WVelocity(); break;
	case NODE_EPressure: EPressure(); break;
	case NODE_NSymmetry: NSymmetry(); break;
	case NODE_SSymmetry: SSymmetry(); break;
	case NODE_ISymmetry: ISymmetry(); break;
	case NODE_Wall:
	case NODE_Solid: BounceBack(); break;
	}
	if ((NodeType & NODE_COLLISION) == NODE_MRT) CollisionMRT();
}""")
    A("CudaDeviceFunction void SetEquilibrum(real_t rho, real_t Jx, real_t Jy, real_t Jz) {")
    A(c_assign(f, EQ.feq))
    A("}")
    A("""CudaDeviceFunction void Init() {
	vector_t ST = SyntheticTurbulence(X,Y,Z);
	real_t rho = 1.0 + Pressure*3.0;
	SetEquilibrum(rho, rho*(Velocity + Turbulence*ST.x), rho*Turbulence*ST.y, rho*Turbulence*ST.z);
}""")
    A("CudaDeviceFunction void CollisionMRT() {")
    _orthogonal_mrt_collision(A, U, f, ["Jx", "Jy", "Jz"], ["ForceX", "ForceY", "ForceZ"], smag="Smag",
                              les_flag="(NodeType & NODE_LES) == NODE_Smagorinsky",
                              extra="\tAddToXFlux(Jx/rho);\n\tAddToYFlux(Jy/rho);\n\tAddToZFlux(Jz/rho);")
    A("}")
    m.code = "\n".join(c)
    return m


MODELS = {
    "d2q9_SRT": ModelSpec("d2q9_SRT", build_d2q9_srt, reference="models/flow/d2q9_SRT"),
    "d3q27": ModelSpec("d3q27", build_d3q27, reference="models/flow/d3q27"),
}
