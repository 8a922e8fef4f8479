"""d3q19 MRT flow model.

Model surface of the reference's ``models/flow/d3q19`` (Dynamics.R:1-95, Dynamics.c.Rt:1-290):
classic 19-moment orthogonal MRT (d'Humieres basis) with two relaxation factors derived from
``omega``, on-site W/E velocity and pressure faces with transverse-momentum correction, full
bounce-back walls, slice and volume Globals and a MAX global.
"""
from __future__ import annotations

import numpy as np

from ..gen.lbmlib import d3q19_mrt_matrix, d3q19_velocities, poly_moment_matrix, sym_moments, sym_populations
from ..gen.modeldef import ModelDef
from ..gen.poly import PV, Poly, PVec, c_assign, psum, to_c

NAME = "d3q19"
OPTIONS: list = []
VARIANTS = [dict()]


def build(options=None) -> ModelDef:
    m = ModelDef(NAME, dict(options or {}))
    U = d3q19_velocities()
    for i, u in enumerate(U):
        m.AddDensity(name=f"f{i}", dx=int(u[0]), dy=int(u[1]), dz=int(u[2]), comment=f"density F {i}", group="f")
    m.AddQuantity(name="P", unit="Pa")
    m.AddQuantity(name="U", unit="m/s", vector=True)
    m.AddSetting(name="omega", comment="One over relaxation time")
    m.AddSetting(name="nu", omega="1.0/(3*nu + 0.5)", default=0.16666666, comment="Viscosity", unit="1m2/s")
    m.AddSetting(name="InletVelocity", default="0m/s", comment="Inlet velocity", unit="1m/s")
    m.AddSetting(name="InletPressure", InletDensity="1.0+InletPressure*3", default="0Pa", comment="Inlet pressure", unit="1Pa")
    m.AddSetting(name="InletDensity", default=1, comment="Inlet density", unit="1kg/m3")
    m.AddSetting(name="ForceX", comment="Force force X")
    m.AddSetting(name="ForceY", comment="Force force Y")
    m.AddSetting(name="ForceZ", comment="Force force Z")
    m.AddGlobal(name="Flux", comment="Volume flux", unit="m3/s")
    for p in ["XY", "XZ", "YZ"]:
        m.AddNodeType(name=f"{p}slice", group="ADDITIONALS")
    for p in ["XY", "XZ", "YZ"]:
        for q, unit in [("vx", "m3/s"), ("vy", "m3/s"), ("vz", "m3/s"), ("rho", "kg/m"), ("area", "m2")]:
            m.AddGlobal(name=f"{p}{q}", comment="Volume flux", unit=unit)
    for q, unit in [("vx", "m4/s"), ("vy", "m4/s"), ("vz", "m4/s"), ("px", "mkg/s"), ("py", "mkg/s"), ("pz", "mkg/s"),
                    ("rho", "kg"), ("volume", "m3")]:
        m.AddGlobal(name=f"VOL{q}", comment="Volume flux", unit=unit)
    m.AddGlobal(name="MaxV", comment="Max velocity", unit="m3", op="MAX")
    for n in ["EPressure", "Solid", "Wall", "WPressure", "WPressureL", "WVelocity"]:
        m.AddNodeType(name=n, group="BOUNDARY")
    m.AddNodeType(name="MRT", group="COLLISION")
    m.code = _physics(m, U)
    return m


def _physics(m: ModelDef, U) -> str:
    f = PV("f", range(19))
    Mm = d3q19_mrt_matrix()                 # (moments x velocities)
    MT = Mm.T                               # f @ MT = moments
    norm = (Mm ** 2).sum(axis=1)
    X, Y, Z = U[:, 0], U[:, 1], U[:, 2]
    rho, J = Poly.sym("rho"), PV(["Jx", "Jy", "Jz"])
    # equilibrium: second order in J, projected on the MRT basis
    B = poly_moment_matrix(U)
    H = []
    sig = Poly.sym("sigma")
    for j in range(19):
        h = rho
        for i in range(3):
            if B.p[j, i] == 1:
                h = h * J[i] * rho ** -1
            elif B.p[j, i] == 2:
                h = h * (J[i] ** 2 * rho ** -2 + sig)
        H.append(h)
    from fractions import Fraction
    feq = (PVec(H) @ np.linalg.inv(B.mat)).subst({"sigma": Fraction(1, 3)})
    Req = PVec([r.truncate(["Jx", "Jy", "Jz"], 2) for r in (feq @ MT)])
    feq2 = Req @ (np.diag(1.0 / norm) @ Mm)
    selU = [3, 5, 7]
    selR = [i for i in range(19) if i not in [0] + selU]
    R = PV("R", range(1, 20))
    Rs = R.copy()
    Rs[0] = rho
    for k, i in enumerate(selU):
        Rs[i] = J[k]
    jx, jy, jz = (f @ U[:, k].astype(float) for k in range(3))
    fs = lambda mask: psum(f[i] for i in np.nonzero(mask)[0])
    c = []
    A = c.append
    A(f"CudaDeviceFunction real_t getRho() {{\n\treturn {to_c(f.sum())};\n}}")
    A(f"CudaDeviceFunction real_t getP() {{\n\treturn (({to_c(f.sum())})-1.0)/3.0;\n}}")
    A("CudaDeviceFunction vector_t getU() {\n\treal_t d = getRho();\n\tvector_t u;")
    A(f"\tu.x = {to_c(jx)};\n\tu.y = {to_c(jy)};\n\tu.z = {to_c(jz)};")
    A("""	u.x /= d;
	u.y /= d;
	u.z /= d;
	if ((NodeType & NODE_BOUNDARY) == NODE_Wall) u.x = u.y = u.z = 0;
	return u;
}
CudaDeviceFunction float2 Color() {
	float2 ret;
	vector_t u = getU();
	ret.x = sqrt(u.x*u.x + u.y*u.y + u.z*u.z);
	ret.y = ((NodeType & NODE_BOUNDARY) == NODE_Wall) ? 0 : 1;
	return ret;
}""")
    A("CudaDeviceFunction void BounceBack() {\n\treal_t uf;")
    done = set()
    for i in range(19):
        for j in range(19):
            if i < j and np.all(U[i] == -U[j]) and (i, j) not in done:
                done.add((i, j))
                A(f"\tuf = f{j};\n\tf{j} = f{i};\n\tf{i} = uf;")
    A("}")
    # on-site face conditions (x-normal), transverse momentum corrections Nxy / Nxz
    nxy = (fs((X == 0) & (Y > 0)) - fs((X == 0) & (Y < 0)))
    nxz = (fs((X == 0) & (Z > 0)) - fs((X == 0) & (Z < 0)))
    idx = {tuple(u): i for i, u in enumerate(U)}

    def inlet(sign):
        s = sign
        out = [f"\tNxy = ({to_c(nxy)})/2.;", f"\tNxz = ({to_c(nxz)})/2.;"]
        pm = "+" if s > 0 else "-"
        mp = "-" if s > 0 else "+"
        out.append(f"\tf{idx[(s,0,0)]} = f{idx[(-s,0,0)]} {pm} rho * ux / 3.;")
        out.append(f"\tf{idx[(s,1,0)]} = f{idx[(-s,-1,0)]} {pm} rho * ux / 6. - Nxy;")
        out.append(f"\tf{idx[(s,-1,0)]} = f{idx[(-s,1,0)]} {pm} rho * ux / 6. + Nxy;")
        out.append(f"\tf{idx[(s,0,1)]} = f{idx[(-s,0,-1)]} {pm} rho * ux / 6. - Nxz;")
        out.append(f"\tf{idx[(s,0,-1)]} = f{idx[(-s,0,1)]} {pm} rho * ux / 6. + Nxz;")
        return "\n".join(out)
    A("CudaDeviceFunction void WVelocity() {\n\treal_t rho, Nxy, Nxz;\n\treal_t ux = InletVelocity;")
    A(f"\trho = {to_c(fs(X < 0))};\n\trho = {to_c(fs(X == 0))} + 2.0*rho;")
    A("\trho = rho / (1. - ux);")
    A(inlet(1) + "\n}")
    A("CudaDeviceFunction void WPressure() {\n\treal_t rho, Nxy, Nxz;\n\treal_t ux;\n\trho = InletDensity;")
    A(f"\tux = {to_c(fs(X < 0))};\n\tux = {to_c(fs(X == 0))} + 2.0*ux;")
    A("\tux = 1. - ux / rho;")
    A(inlet(1) + "\n}")
    A("CudaDeviceFunction void WPressureLimited() {\n\treal_t rho, Nxy, Nxz, SF, ux;")
    A(f"\tSF = {to_c(fs(X < 0))};\n\tSF = {to_c(fs(X == 0))} + 2.0*SF;")
    A("""	rho = InletDensity;
	ux = 1. - SF / rho;
	if (ux > InletVelocity) {
		ux = InletVelocity;
		rho = SF / (1. - ux);
	}""")
    A(inlet(1) + "\n}")
    A("CudaDeviceFunction void EPressure() {\n\treal_t rho = 1.0;\n\treal_t Nxy, Nxz;\n\treal_t ux;")
    A(f"\tux = {to_c(fs(X > 0))};\n\tux = {to_c(fs(X == 0))} + 2.0*ux;")
    A("\tux = -1. + ux / rho;")
    A(inlet(-1) + "\n}")
    A("""CudaDeviceFunction void RunBoundaries() {
	// persistent kernel only, as in d3q27_cumulant: most nodes carry no boundary bits (26 Globals every iteration: 37.0 -> 37.4 GLUPS)
	if (TCLB_NODE_PERSISTENT && (NodeType & NODE_BOUNDARY) == 0) return;
	switch (NodeType & NODE_BOUNDARY) {
	case NODE_WPressureL: WPressureLimited(); break;
	case NODE_WPressure: WPressure(); break;
	case NODE_WVelocity: WVelocity(); break;
	case NODE_EPressure: EPressure(); break;
	case NODE_Wall:
	case NODE_Solid: BounceBack(); break;
	}
}
CudaDeviceFunction void Run() {
	RunBoundaries();
	switch (NodeType & NODE_COLLISION) {
	case NODE_MRT: CollisionMRT(); break;
	}
}""")
    A("CudaDeviceFunction void SetEquilibrum(real_t rho, real_t Jx, real_t Jy, real_t Jz) {")
    A(c_assign(f, feq2))
    A("}")
    A("CudaDeviceFunction void Init() {\n\tSetEquilibrum(1.0, InletVelocity, 0., 0.);\n}")
    A("CudaDeviceFunction void CollisionMRT() {")
    A("\treal_t " + ", ".join(to_c(x) for x in Rs) + ";")
    A("\treal_t gamma1 = 1 - omega;")
    A("\treal_t gamma2 = 1 - 8.0 * ( 2.0 - omega )/( 8.0 - omega );")
    S = [Poly.const(0)] * 19
    for i in [1, 2] + list(range(9, 16)):
        S[i] = Poly.sym("gamma1")
    for i in [4, 6, 8, 16, 17, 18]:
        S[i] = Poly.sym("gamma2")
    S = PVec(S)
    A(sym_moments(f, MT, U, Rs))
    A(c_assign(Rs[selR], (S * (Rs - Req))[selR]))
    A("\tJx += rho*ForceX;\n\tJy += rho*ForceY;\n\tJz += rho*ForceZ;")
    A(c_assign(Rs[selR], (Rs + Req)[selR]))
    A(sym_populations(Rs, np.diag(1.0 / norm) @ Mm, U, f))
    # the reference divides ALL moments by the squared norms of the MRT rows before it forms the velocity for the
    # Globals (`C(R, Mw, eq="/=")`, models/flow/d3q19/Dynamics.c.Rt:263-265), so VOLrho, VOLp*, VOLv*, the slice sums and
    # MaxV are sums of those scaled moments; a log written by this model has to carry the same numbers
    A("\t" + " ".join(f"{to_c(Rs[i])} = {to_c(Rs[i])} * {1.0 / float(norm[i])!r};" for i in range(19)
                     if to_c(Rs[i]) in ("rho", "Jx", "Jy", "Jz")))
    A("""	real_t ux, uy, uz;
	ux = Jx / rho;
	uy = Jy / rho;
	uz = Jz / rho;
	switch (NodeType & NODE_ADDITIONALS) {""")
    for p in ["XY", "XZ", "YZ"]:
        A(f"\tcase NODE_{p}slice:\n\t\tAddTo{p}vx(ux);\n\t\tAddTo{p}vy(uy);\n\t\tAddTo{p}vz(uz);\n\t\tAddTo{p}rho(rho);\n\t\tAddTo{p}area(1);\n\t\tbreak;")
    A("""	}
	AddToVOLvx(ux);
	AddToVOLvy(uy);
	AddToVOLvz(uz);
	AddToVOLpx(Jx);
	AddToVOLpy(Jy);
	AddToVOLpz(Jz);
	AddToVOLrho(rho);
	AddToVOLvolume(1);
	AddToMaxV(sqrt(ux*ux+uy*uy+uz*uz));
}""")
    return "\n".join(c)
