"""d2q9 MRT flow model (options: bc).

Same model surface as the reference's ``models/flow/d2q9`` (Dynamics.R:1-85 declares the
densities/settings/globals/node types; Dynamics.c.Rt:1-269 the physics): MRT collision in a
weight-orthogonal moment basis with one relaxation factor per moment order (S2,S3,S4),
Zou-He velocity/pressure closures on W/E/N/S faces, symmetry, full bounce-back, pressure-loss
and flux Globals on Inlet/Outlet objective nodes.  Written against this framework's own
generator library (:mod:`tclb_b200.gen.lbmlib`).
"""
from __future__ import annotations

import numpy as np

from ..gen.lbmlib import (d2q9_velocities, full_bounce_back, lattice_weights, mrt_eq, sym_moments, sym_populations, symmetry,
                          weighted_cholesky_basis, zouhe_rewrite)
from ..gen.modeldef import ModelDef
from ..gen.poly import PV, Poly, PVec, c_assign, to_c

NAME = "d2q9"
OPTIONS = ["bc", "autosym", "autosym2"]      # the reference's OPT="bc*autosym" (variants autosym = 1, autosym2 = 2: src/models.R:52-57)
VARIANTS = [dict(), dict(bc=True), dict(autosym=True), dict(autosym2=True), dict(bc=True, autosym=True), dict(bc=True, autosym2=True)]


def build(options=None) -> ModelDef:
    opt = dict(bc=False)
    opt.update(options or {})
    sym = 2 if opt.pop("autosym2", False) else (1 if opt.get("autosym") else 0)
    opt["autosym"] = sym
    m = ModelDef(NAME, opt)
    U = d2q9_velocities()
    w = [4 / 9] + [1 / 9] * 4 + [1 / 36] * 4
    for i, (u, wi) in enumerate(zip(U, w)):
        m.AddDensity(name=f"f[{i}]", dx=int(u[0]), dy=int(u[1]), group="f", shift=wi)
    if opt["bc"]:
        m.AddDensity(name="BC[0]", group="BC", parameter=True)
        m.AddDensity(name="BC[1]", group="BC", parameter=True)
    m.AddQuantity(name="Rho", unit="kg/m3")
    m.AddQuantity(name="U", unit="m/s", vector=True)
    m.AddSetting(name="RelaxationRate", S2="1-RelaxationRate", comment="one over relaxation time")
    m.AddSetting(name="Viscosity", RelaxationRate="1.0/(3*Viscosity + 0.5)", default=0.16666666, comment="viscosity")
    m.AddSetting(name="VelocityX", default=0, comment="inlet/outlet/init velocity", zonal=True)
    m.AddSetting(name="VelocityY", default=0, comment="inlet/outlet/init velocity", zonal=True)
    m.AddSetting(name="Pressure", default=0, comment="inlet/outlet/init density", zonal=True)
    m.AddSetting(name="GravitationX")
    m.AddSetting(name="GravitationY")
    m.AddGlobal(name="PressureLoss", comment="pressure loss", unit="1mPa")
    m.AddGlobal(name="OutletFlux", comment="pressure loss", unit="1m2/s")
    m.AddGlobal(name="InletFlux", comment="pressure loss", unit="1m2/s")
    m.AddSetting(name="S2", default="0", comment="MRT Sx")
    m.AddSetting(name="S3", default="0", comment="MRT Sx")
    m.AddSetting(name="S4", default="0", comment="MRT Sx")
    for n in ["EPressure", "WPressure", "NVelocity", "SVelocity", "WVelocity", "EVelocity", "NSymmetry", "SSymmetry"]:
        m.AddNodeType(name=n, group="BOUNDARY")
    m.AddNodeType(name="Inlet", group="OBJECTIVE")
    m.AddNodeType(name="Outlet", group="OBJECTIVE")
    m.AddNodeType(name="Solid", group="BOUNDARY")
    m.AddNodeType(name="Wall", group="BOUNDARY")
    m.AddNodeType(name="MRT", group="COLLISION")
    m.code = _physics(m, U, opt)
    return m


def _physics(m: ModelDef, U: np.ndarray, opt) -> str:
    f = PV("f[", range(9), "]")
    rho = Poly.sym("rho")
    J = PV(["Jx", "Jy"])
    dens = m.density_dicts()
    wi = lattice_weights(U)
    s3, s2, s6 = np.sqrt(3.0), np.sqrt(2.0), np.sqrt(6.0)
    B = weighted_cholesky_basis(U, [1, s3 / 3, s3 / 3, s2, s2, 1, s6 / 3, s6 / 3, 2])
    EQ = mrt_eq(U, mat=B)
    Binv = np.linalg.inv(B)
    bcv = PV(["BC[0]", "BC[1]"]) if opt["bc"] else PV(["VelocityX", "VelocityY"])
    Ux = PVec([Poly.const(int(u[0])) for u in U])
    Uy = PVec([Poly.const(int(u[1])) for u in U])
    jx = f @ U[:, 0].astype(float)
    jy = f @ U[:, 1].astype(float)
    c = []
    A = c.append
    A("#define pi 3.141592653589793116")
    A(f"CudaDeviceFunction real_t getRho() {{\n\treturn {to_c(f.sum())};\n}}")
    A("CudaDeviceFunction vector_t getU() {")
    A(f"\treal_t d = {to_c(f.sum())};")
    A("\tvector_t u;")
    A(f"\tu.x = {to_c(jx)};\n\tu.y = {to_c(jy)};")
    A("\tu.x /= d;\n\tu.y /= d;")
    if opt["bc"]:
        A("\tif (!IamBOUNDARY) {\n\t\tu.x += BC[0]*0.5;\n\t\tu.y += BC[1]*0.5;\n\t}")
    A("\tu.x += GravitationX*0.5;\n\tu.y += GravitationY*0.5;\n\tu.z = 0.0;\n\treturn u;\n}")
    A("""CudaDeviceFunction float2 Color() {
	float2 ret;
	vector_t u = getU();
	ret.x = sqrt(u.x*u.x + u.y*u.y);
	ret.y = (NodeType == NODE_Solid) ? 0 : 1;
	return ret;
}""")
    A("CudaDeviceFunction void SetEquilibrum(real_t rho, real_t Jx, real_t Jy) {")
    A(c_assign(f, EQ.feq))
    if opt["bc"]:
        A("\tif (IamBOUNDARY) {\n\t\tBC[0] = Jx / rho;\n\t\tBC[1] = Jy / rho;\n\t} else {\n\t\tBC[0] = 0;\n\t\tBC[1] = 0;\n\t}")
    A("}")
    A("""CudaDeviceFunction void Init() {
	real_t rho, ux, uy;
	rho = (1+Pressure*3);
	ux = VelocityX;
	uy = VelocityY;
	if (IamWall) {
		rho = 1;
		ux = 0;
		uy = 0;
	}
	SetEquilibrum(rho, ux*rho, uy*rho);
}""")
    A("""CudaDeviceFunction void Run() {
	switch (NodeType & NODE_BOUNDARY) {
	case NODE_Solid:
	case NODE_Wall: BounceBack(); break;
	case NODE_EVelocity: EVelocity(); break;
	case NODE_WPressure: WPressure(); break;
	case NODE_WVelocity: WVelocity(); break;
	case NODE_EPressure: EPressure(); break;
	case NODE_NVelocity: NVelocity(); break;
	case NODE_SVelocity: SVelocity(); break;
	case NODE_NSymmetry: NSymmetry(); break;
	case NODE_SSymmetry: SSymmetry(); break;
	}
	if (NodeType & NODE_MRT) CollisionMRT();
}""")
    A("CudaDeviceFunction void BounceBack() {")
    if opt["bc"]:
        A(f"\treal_t temp = {to_c(f.sum())};")
        A("\tif ( temp != 1+Pressure*3 ) {")
        A(f"\t\tBC[0] = {to_c(jx * 2)};\n\t\tBC[1] = {to_c(jy * 2)};")
        A("\t} else {\n\t\tBC[0] = 0.0;\n\t\tBC[1] = 0.0;\n\t}")
    A(full_bounce_back(dens, "f"))
    A("}")
    for nm, n, kind, rhs in [("EVelocity", (1, 0), "velocity", bcv), ("WPressure", (-1, 0), "pressure", None),
                             ("WVelocity", (-1, 0), "velocity", bcv), ("EPressure", (1, 0), "pressure", None),
                             ("NVelocity", (0, 1), "velocity", bcv), ("SVelocity", (0, -1), "velocity", bcv)]:
        A(f"CudaDeviceFunction void {nm}() {{")
        A(zouhe_rewrite(U, f, n, kind, rhs))
        A("}")
    A("CudaDeviceFunction void NSymmetry() {\n" + symmetry(dens, 2, -1, "f") + "}")
    A("CudaDeviceFunction void SSymmetry() {\n" + symmetry(dens, 2, 1, "f") + "}")

    # ---- collision ----
    R = PV("R", range(9))
    selR = EQ.order > 1
    Rn = R.copy()
    for i in np.nonzero(~selR)[0]:
        Rn[i] = EQ.Req[i]                      # rho, Jx, Jy stand for the conserved moments
    S = PVec([Poly.sym("S2") if o == 2 else Poly.sym("S3") if o == 3 else Poly.sym("S4") if o == 4 else Poly.const(0)
              for o in EQ.order])
    names = ", ".join(to_c(x) for x in Rn)
    A("CudaDeviceFunction void CollisionMRT() {")
    A(f"\treal_t {names};")
    A("\treal_t Usq = 0;")
    A(sym_moments(f, EQ.mat, U, Rn))
    usq = (J @ J) * rho ** -2
    A("\tswitch (NodeType & NODE_OBJECTIVE) {")
    A("\tcase NODE_Outlet:")
    A(f"\t\tUsq = {to_c(usq)};")
    A("\t\tAddToOutletFlux(Jx/rho/rho);")
    A("\t\tAddToPressureLoss(-(Jx/rho)/rho*((rho-1.)/3. + Usq/rho/2.));")
    A("\t\tbreak;")
    A("\tcase NODE_Inlet:")
    A(f"\t\tUsq = {to_c(usq)};")
    A("\t\tAddToInletFlux(Jx/rho/rho);")
    A("\t\tAddToPressureLoss((Jx/rho)/rho*((rho-1.)/3. + Usq/rho/2.));")
    A("\t\tbreak;")
    A("\t}")
    sel = np.nonzero(selR)[0]
    A(c_assign(Rn[sel], (Rn - EQ.Req)[sel]))
    A(c_assign(Rn[sel], (Rn * S)[sel]))
    if opt["bc"]:
        A("\tif (!IamBOUNDARY) {\n\t\tJx = Jx + (GravitationX + BC[0])*rho;\n\t\tJy = Jy + (GravitationY + BC[1])*rho;\n\t}")
    else:
        A("\tJx = Jx + GravitationX*rho;\n\tJy = Jy + GravitationY*rho;")
    A(c_assign(Rn[sel], (Rn + EQ.Req)[sel]))
    A(sym_populations(Rn, Binv, U, f))
    A("}")
    return "\n".join(c)
