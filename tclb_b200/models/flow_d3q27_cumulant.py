"""d3q27 cumulant flow model (options AVG, IB, SMAG).

Model surface of the reference's ``models/flow/d3q27_cumulant`` (Dynamics.R:1-98: 27 densities
f000..f222 plus same-node field access, settings nu/nubuffer/Velocity/Pressure/Turbulence/
GalileanCorrection/Force*/Omega[/Smag], Globals Density/Flux/Drag/Lift/Lateral/Mass/*Momentum,
node types; Dynamics.c.Rt:1-520: raw-moment sweeps, cumulant recursion, relaxation with Galilean
correction, Zou-He faces, symmetry, interpolated bounce-back on wall cuts, running averages).
The transforms are generated from the combinatorial recursion in :mod:`tclb_b200.gen.lbmlib`.
"""
from __future__ import annotations

import numpy as np

from ..gen.lbmlib import (bounce_index, central_moment_backward, central_moment_forward, central_moment_forward_shared, cumulant_from_moments, full_bounce_back, moments_from_cumulant, mrt_eq,
                          raw_moment_sweeps, symmetry, tensor_names, tensor_velocities, zouhe)
from ..gen.modeldef import ModelDef
from ..gen.poly import PV, Poly, PVec, c_assign, to_c

NAME = "d3q27_cumulant"
OPTIONS = ["AVG", "IB", "SMAG"]
VARIANTS = [dict(), dict(AVG=True), dict(IB=True), dict(SMAG=True), dict(AVG=True, IB=True), dict(AVG=True, SMAG=True),
            dict(IB=True, SMAG=True), dict(AVG=True, IB=True, SMAG=True)]


def build(options=None, collision="central") -> ModelDef:
    """`collision`: "central" = central-moment sweeps + closed-form 4th..6th order closure (default, ~40% of the
    floating-point work); "raw" = raw moments + generic cumulant recursion (the reference's arrangement, kept to
    test that both give the same result)."""
    opt = dict(AVG=False, IB=False, SMAG=False)
    opt.update(options or {})
    m = ModelDef(NAME, opt)
    U = tensor_velocities(3)
    fname = tensor_names("f", 3)
    for i, (n, u) in enumerate(zip(fname, U)):
        m.AddDensity(name=n, dx=int(u[0]), dy=int(u[1]), dz=int(u[2]), comment=f"density F {i}", group="f")
    m.AddQuantity(name="P", unit="Pa")
    m.AddQuantity(name="U", unit="m/s", vector=True)
    m.AddQuantity(name="Solid", unit="1")
    m.AddSetting(name="nu", default=0.16666666, comment="Viscosity")
    m.AddSetting(name="nubuffer", default=0.01, comment="Viscosity in the buffer layer")
    m.AddSetting(name="Velocity", default="0m/s", comment="Inlet velocity", zonal=True)
    m.AddSetting(name="Pressure", default="0Pa", comment="Inlet pressure", zonal=True)
    m.AddSetting(name="Turbulence", comment="Turbulence intensity", zonal=True)
    m.AddSetting(name="GalileanCorrection", default=1.0, comment="Galilean correction term")
    m.AddSetting(name="ForceX", default=0, comment="Force force X")
    m.AddSetting(name="ForceY", default=0, comment="Force force Y")
    m.AddSetting(name="ForceZ", default=0, comment="Force force Z")
    m.AddSetting(name="Omega", default=1, comment="relaxation rate for 3rd order cumulants")
    m.AddGlobal(name="Density", comment="system density", unit="kg/m3")
    m.AddGlobal(name="Flux", comment="Volume flux", unit="m3/s")
    m.AddGlobal(name="Drag", comment="Force exerted on body in X-direction", unit="N")
    m.AddGlobal(name="Lift", comment="Force exerted on body in Z-direction", unit="N")
    m.AddGlobal(name="Lateral", comment="Force exerted on body in Y-direction", unit="N")
    m.AddGlobal(name="Mass", comment="Integral of density over the domain", unit="kg")
    m.AddGlobal(name="XMomentum", comment="Integral of momentum in X", unit="kgm/s")
    m.AddGlobal(name="YMomentum", comment="Integral of momentum in Y", unit="kgm/s")
    m.AddGlobal(name="ZMomentum", comment="Integral of momentum in Z", unit="kgm/s")
    for n in ["Buffer", "WVelocityTurbulent", "NVelocity", "SVelocity", "NPressure", "SPressure"]:
        m.AddNodeType(name=n, group="BOUNDARY")
    m.AddNodeType(name="NSymmetry", group="ADDITIONALS")
    m.AddNodeType(name="SSymmetry", group="ADDITIONALS")
    m.AddNodeType(name="Body", group="BODY")
    for n in fname:
        m.AddField(n, dx=0, dy=0, dz=0)          # same-node access to the pre-streaming value
    if opt["SMAG"]:
        m.AddSetting(name="Smag", default=0, comment="Smagorinsky coefficient for SGS modeling")
    if opt["IB"]:
        m.AddNodeType(name="IB", group="HO_BOUNDARY")
    if opt["AVG"]:
        m.AddQuantity(name="KinE", comment="Turbulent kinetic energy")
        m.AddQuantity(name="ReStr", comment="Reynolds stress off-diagonal component", vector=True)
        m.AddQuantity(name="Dissipation", comment="Dissipation e")
        m.AddQuantity(name="avgU", unit="m/s", vector=True)
        m.AddQuantity(name="varU", vector=True)
        m.AddQuantity(name="averageP", unit="Pa")
        for n in ["avgP", "varUX", "varUY", "varUZ", "varUXUY", "varUXUZ", "varUYUZ", "avgdxu2", "avgdyv2", "avgdzw2"]:
            m.AddDensity(name=n, dx=0, dy=0, dz=0, average=True)
        for n in ["avgUX", "avgUY", "avgUZ"]:
            m.AddDensity(name=n, average=True)
        m.AddField(name="avgUX", dx=[-1, 1], average=True)
        m.AddField(name="avgUY", dy=[-1, 1], average=True)
        m.AddField(name="avgUZ", dz=[1, -1], average=True)
    for n in ["EPressure", "EVelocity", "Solid", "Wall", "WPressure", "WVelocity"]:
        m.AddNodeType(name=n, group="BOUNDARY")
    m.AddNodeType(name="MRT", group="COLLISION")
    m.code = _physics(m, U, fname, opt, collision)
    # measured on B200 (profiles/): 64 registers/thread -> 8 blocks of 128 threads per SM is the fastest fp32 build
    # measured on B200 (1024x512x512 fp32): plain kernel 64 regs/8 blocks 28.8 GLUPS (72 regs: 24.9, 80: 23.9);
    # Globals kernel 80 regs/6 blocks 24.3 GLUPS (64 regs: 20.8-23.3, 72: 22.4, 96: 18.9)
    # prec_div: IEEE fp32 division kept (measured: one-block-per-tile kernel 28.0 GLUPS with it, 26.0 without; ring 26.5 / 26.8)
    m.tuning = dict(min_blocks_float=8, min_blocks_double=3, min_blocks_g_float=6, prec_div=True)
    return m


def _physics(m: ModelDef, U, fname, opt, collision="central") -> str:
    f = PV(fname)
    dens = m.density_dicts("f")
    EQ = mrt_eq(U, "rho", ("Jx", "Jy", "Jz"), orthogonal=False)
    jx, jy, jz = (f @ U[:, k].astype(float) for k in range(3))
    c = []
    A = c.append
    A(f"CudaDeviceFunction real_t getRho() {{\n\treturn {to_c(f.sum())};\n}}")
    A(f"CudaDeviceFunction real_t getP() {{\n\treturn (({to_c(f.sum())})-1.0)/3.0;\n}}")
    A("CudaDeviceFunction vector_t getU_() {")
    A("\treal_t d = getRho();\n\tvector_t u;")
    A(f"\tu.x = {to_c(jx)};\n\tu.y = {to_c(jy)};\n\tu.z = {to_c(jz)};")
    A("""	if ((NodeType & NODE_BOUNDARY) == NODE_Wall) {
		u.x = 0; u.y = 0; u.z = 0;
	} else {
		u.x = (u.x + ForceX/2.)/d;
		u.y = (u.y + ForceY/2.)/d;
		u.z = (u.z + ForceZ/2.)/d;
	}
	return u;
}
CudaDeviceFunction vector_t getU() {
	RunBoundaries();
	return getU_();
}
CudaDeviceFunction real_t getSolid() {
	return ((NodeType & NODE_BOUNDARY) == NODE_Wall) ? 1 : 0;
}""")
    if opt["AVG"]:
        A("""CudaDeviceFunction real_t getaverageP() { return avgP/average_iter; }
CudaDeviceFunction real_t getDissipation() {
	real_t ai = average_iter;
	real_t gx = avgUX(1,0,0) - avgUX(-1,0,0), gy = avgUY(0,1,0) - avgUY(0,-1,0), gz = avgUZ(0,0,1) - avgUZ(0,0,-1);
	real_t diss = nu*((avgdxu2/ai) - gx*gx/(4*ai*ai));
	diss = diss + nu*((avgdyv2/ai) - gy*gy/(4*ai*ai));
	diss = diss + nu*((avgdzw2/ai) - gz*gz/(4*ai*ai));
	return diss;
}
CudaDeviceFunction vector_t getavgU() {
	vector_t u;
	u.x = avgUX(0,0,0)/average_iter;
	u.y = avgUY(0,0,0)/average_iter;
	u.z = avgUZ(0,0,0)/average_iter;
	return u;
}
CudaDeviceFunction vector_t getvarU() {
	vector_t u, a = getavgU();
	u.x = varUX/average_iter - a.x*a.x;
	u.y = varUY/average_iter - a.y*a.y;
	u.z = varUZ/average_iter - a.z*a.z;
	return u;
}
CudaDeviceFunction vector_t getReStr() {
	vector_t st, a = getavgU();
	st.x = varUYUZ/average_iter - a.y*a.z;
	st.y = varUXUZ/average_iter - a.x*a.z;
	st.z = varUXUY/average_iter - a.x*a.y;
	return st;
}
CudaDeviceFunction real_t getKinE() {
	vector_t v = getvarU();
	return 0.5*(v.x + v.y + v.z);
}""")
    A("""CudaDeviceFunction float2 Color() {
	float2 ret;
	vector_t u = getU();
	ret.x = sqrt(u.x*u.x + u.y*u.y + u.z*u.z);
	ret.y = (NodeType == NODE_Solid) ? 0 : 1;
	return ret;
}""")
    A("CudaDeviceFunction void BounceBack() {\n\tvector_t p1, p2;")
    A(f"\tp1.x = {to_c(jx)};\n\tp1.y = {to_c(jy)};\n\tp1.z = {to_c(jz)};")
    A(full_bounce_back(dens, "f"))
    A(f"\tp2.x = {to_c(jx)};\n\tp2.y = {to_c(jy)};\n\tp2.z = {to_c(jz)};")
    A("""	if ((NodeType & NODE_BODY) == NODE_Body) {
		AddToDrag(-(p2.x-p1.x));
		AddToLateral(-(p2.y-p1.y));
		AddToLift(-(p2.z-p1.z));
	}
}""")
    for nm, direction, sign, kind in [("EVelocity", 1, -1, "velocity"), ("WVelocity", 1, 1, "velocity"),
                                      ("SVelocity", 2, 1, "velocity"), ("NVelocity", 2, -1, "velocity"),
                                      ("WPressure", 1, 1, "pressure"), ("SPressure", 2, 1, "pressure"),
                                      ("NPressure", 2, -1, "pressure"), ("EPressure", 1, -1, "pressure")]:
        A(f"CudaDeviceFunction void {nm}() {{")
        A(zouhe(EQ, f, direction, sign, kind))
        A("}")
    A("CudaDeviceFunction void NSymmetry() {\n" + symmetry(dens, 2, -1) + "}")
    A("CudaDeviceFunction void SSymmetry() {\n" + symmetry(dens, 2, 1) + "}")
    if opt["IB"]:
        b = bounce_index(U)
        A("CudaDeviceFunction void InterpolatedBounceBack() {\n\treal_t q;\n\tvector_t F;\n\tF.x = F.y = F.z = 0.0;")
        qnames = tensor_names("Q", 3)
        for i in range(1, 27):
            fo, fb = fname[i], fname[b[i]]
            A(f"\tq = {qnames[i]};")
            A("\tif (q != NO_CUT) {")
            A("\t\treal_t p = ((real_t)q) * (2.0/CUT_MAX);")
            A(f"\t\treal_t fa = {fo}(0,0,0);")
            A(f"\t\tif (p > 1.0) {{\n\t\t\t{fb} = fa/p + {fb}(0,0,0)*(p-1)/p;\n\t\t}} else {{\n\t\t\t{fb} = fa*p + {fo}*(1-p);\n\t\t}}")
            for k, ax in enumerate("xyz"):
                if U[i, k] != 0:
                    sgn = "+" if U[i, k] > 0 else "-"
                    A(f"\t\tF.{ax} = F.{ax} {sgn} ({fb} - fa);")
            A("\t}")
        A("""	if ((NodeType & NODE_BODY) == NODE_Body) {
		AddToDrag(F.x);
		AddToLateral(F.y);
		AddToLift(F.z);
	}
}""")
    A("CudaDeviceFunction void RunBoundaries() {")
    # Persistent kernel only (measured: ring +6 %, one-block-per-tile kernel -22 % with the same line): most nodes carry no
    # boundary bits, leaving before the dispatch keeps the arms' register shuffles off the common path
    A("\tif (TCLB_NODE_PERSISTENT && (NodeType & (NODE_ADDITIONALS | NODE_BOUNDARY" + (" | NODE_HO_BOUNDARY" if opt["IB"] else "") + ")) == 0) return;")
    if opt["IB"]:
        A("\tif ((NodeType & NODE_HO_BOUNDARY) == NODE_IB) InterpolatedBounceBack();")
    A("""	switch (NodeType & NODE_ADDITIONALS) {
	case NODE_NSymmetry: NSymmetry(); break;
	case NODE_SSymmetry: SSymmetry(); break;
	}
	switch (NodeType & NODE_BOUNDARY) {
	case NODE_WVelocityTurbulent: WVelocityTurbulent(); break;
	case NODE_EPressure: EPressure(); break;
	case NODE_WPressure: WPressure(); break;
	case NODE_SPressure: SPressure(); break;
	case NODE_NPressure: NPressure(); break;
	case NODE_WVelocity: WVelocity(); break;
	case NODE_NVelocity: NVelocity(); break;
	case NODE_SVelocity: SVelocity(); break;
	case NODE_EVelocity: EVelocity(); break;
	case NODE_Wall: BounceBack(); break;
	}
}
CudaDeviceFunction void Run() {
	RunBoundaries();
	switch (NodeType & NODE_COLLISION) {
	case NODE_MRT: CollisionMRT(); break;
	}
}""")
    A("CudaDeviceFunction void SetEquilibrum(real_t rho, real_t Jx, real_t Jy, real_t Jz) {")
    A(c_assign(f, EQ.feq))
    A("}")
    A("""CudaDeviceFunction void Init() {
	vector_t ST = SyntheticTurbulence(X,Y,Z);
	ST.x = Turbulence*ST.x + Velocity;
	ST.y = Turbulence*ST.y;
	ST.z = Turbulence*ST.z;
	if ((NodeType & NODE_BOUNDARY) == NODE_Wall) SetEquilibrum(1.0, 0, 0, 0);
	else SetEquilibrum(1.0 + Pressure*3.0, ST.x, ST.y, ST.z);""")
    A("}")
    A("""CudaDeviceFunction void WVelocityTurbulent() {
	vector_t ST = SyntheticTurbulence(X-Time*Velocity,Y,Z);
	ST.x = Turbulence*ST.x + Velocity;
	ST.y = Turbulence*ST.y;
	ST.z = Turbulence*ST.z;""")
    A(zouhe(EQ, f, 1, 1, "velocity", V="ST.x", V3=["ST.x", "ST.y", "ST.z"]))
    A("}")
    # ---- collision ----
    if collision == "raw":
        _collision_raw(A, opt)
    else:
        _collision_central(A, opt, to_c(jx), to_c(jy), to_c(jz))
    return "\n".join(c)


_COLLISION_PROLOGUE = """	real_t rho = getRho();
	vector_t velocity = getU_();
	AddToDensity( rho );
	AddToFlux( velocity.x );
	AddToMass( rho );
	AddToXMomentum( rho * velocity.x );
	AddToYMomentum( rho * velocity.y );
	AddToZMomentum( rho * velocity.z );
	real_t w0 = 1.0/(3*nu+0.5), w1 = 1.0;
	if ((NodeType & NODE_BOUNDARY) != 0) w0 = 1.0/(3*nubuffer+0.5);"""

_SMAG = """	{
		real_t tau_0 = (3*nu + 0.5);
		real_t qq = sqrt((c200-1.0/3)*(c200-1.0/3)+(c020-1.0/3)*(c020-1.0/3)+(c002-1.0/3)*(c002-1.0/3)+2*(c110*c110)+2*(c101*c101)+2*(c011*c011));
		real_t tau_t = 0.5*(sqrt(tau_0*tau_0 + 18*Smag*Smag*qq) - tau_0);
		w0 = 1.0/(tau_0 + tau_t);
		if ((NodeType & NODE_BOUNDARY) != 0) w0 = 1.0/(3*nubuffer+0.5);
	}"""

_RELAX_2ND = """	real_t dxu, dyv, dzw;
	dxu = - w0/(2.)*(2*c200 - c020 - c002) - w1/(2.)*(c200 + c020 + c002 - 1.);
	dyv = dxu + 3.*w0/2.*(c200 - c020);
	dzw = dxu + 3.*w0/2.*(c200 - c002);
%(avg)s	real_t gcor1 = 3.*(1 - w0/2.)*(u.x*u.x*dxu - u.y*u.y*dyv);
	real_t gcor2 = 3.*(1 - w0/2.)*(u.x*u.x*dxu - u.z*u.z*dzw);
	real_t gcor3 = 3.*(1 - w1/2.)*(u.x*u.x*dxu + u.y*u.y*dyv + u.z*u.z*dzw);
	real_t ra, rb, rc;
	ra = (1 - w0)*(c200 - c020);
	rb = (1 - w0)*(c200 - c002);
	rc = w1 + (1 - w1)*(c200 + c020 + c002);
	ra = ra - gcor1 * GalileanCorrection;
	rb = rb - gcor2 * GalileanCorrection;
	rc = rc - gcor3 * GalileanCorrection;
	c200 = (ra + rb + rc)/3.;
	c020 = (rc - 2*ra + rb)/3.;
	c002 = (rc - 2*rb + ra)/3.;
	c110 = c110 * (1-w0);
	c011 = c011 * (1-w0);
	c101 = c101 * (1-w0);"""

_AVG_TAIL = """	{
		vector_t uu = getU_();
		avgP = avgP + getP();
		avgUX = avgUX(0,0,0) + uu.x;
		avgUY = avgUY(0,0,0) + uu.y;
		avgUZ = avgUZ(0,0,0) + uu.z;
		varUX = varUX + uu.x*uu.x;
		varUY = varUY + uu.y*uu.y;
		varUZ = varUZ + uu.z*uu.z;
		varUXUY = varUXUY + uu.x*uu.y;
		varUXUZ = varUXUZ + uu.x*uu.z;
		varUYUZ = varUYUZ + uu.y*uu.z;
	}"""


def _collision_raw(A, opt):
    """raw moments -> cumulants by the generic recursion (the reference's arrangement)."""
    cum = tensor_names("c", 3)
    P = np.array([[int(ch) for ch in n[1:]] for n in cum])
    A("CudaDeviceFunction void CollisionMRT() {")
    A(_COLLISION_PROLOGUE)
    A(raw_moment_sweeps("f", forward=True))
    A("\treal_t " + ", ".join(cum[1:]) + ";")
    A(cumulant_from_moments("f", "c"))
    A("""	vector_t u;
	u.x = c100 + ForceX/(2.*f000);
	u.y = c010 + ForceY/(2.*f000);
	u.z = c001 + ForceZ/(2.*f000);""")
    if opt["SMAG"]:
        A(_SMAG)
    avg = "\tavgdxu2 = avgdxu2 + dxu*dxu;\n\tavgdyv2 = avgdyv2 + dyv*dyv;\n\tavgdzw2 = avgdzw2 + dzw*dzw;\n" if opt["AVG"] else ""
    A(_RELAX_2ND % dict(avg=avg))
    A("\tc100 = c100 + ForceX;\n\tc010 = c010 + ForceY;\n\tc001 = c001 + ForceZ;")
    for n, p in zip(cum, P):
        if p.sum() == 3:
            A(f"\t{n} = (1-Omega)*{n};")
    for n, p in zip(cum, P):
        if p.sum() > 3:
            A(f"\t{n} = 0.0;")
    A(moments_from_cumulant("f", "c"))
    A(raw_moment_sweeps("f", forward=False))
    if opt["AVG"]:
        A(_AVG_TAIL)
    A("}")


def _collision_central(A, opt, jx="0", jy="0", jz="0"):
    """Same collision operator, arranged for the FP32 issue rate of the SM: central-moment sweeps about the
    node velocity (6 FMA-class ops per triple), cumulants of order 2 and 3 are the central moments themselves,
    and because every cumulant of order >= 4 relaxes to zero the post-collision central moments of order 4..6
    follow in closed form from the relaxed 2nd/3rd-order ones - the pre-collision high-order cumulants are never
    formed.  The first-order shift by the body force enters as the velocity of the backward sweep."""
    A("CudaDeviceFunction void CollisionMRT() {")
    A("	vector_t u0;      // momentum/density without the half-force shift: the first raw cumulant")
    # the Globals only need density and velocity: they are accumulated right after those exist, before the sweeps
    # (the point of highest register pressure comes later); momentum is rho*velocity = J + F/2 off the walls
    globals_block = """	{
		const bool wall_ = (NodeType & NODE_BOUNDARY) == NODE_Wall;
		AddToDensity( rho );
		AddToMass( rho );
		AddToFlux( wall_ ? (real_t)0 : u0.x + ForceX*0.5*irho );
		AddToXMomentum( wall_ ? (real_t)0 : jx + ForceX*0.5 );
		AddToYMomentum( wall_ ? (real_t)0 : jy + ForceY*0.5 );
		AddToZMomentum( wall_ ? (real_t)0 : jz + ForceZ*0.5 );
	}"""
    A(central_moment_forward_shared("f", "rho", ["jx", "jy", "jz"], ["u0.x", "u0.y", "u0.z"], "irho", after_velocity=globals_block))
    A("""	real_t w0 = 1.0/(3*nu+0.5), w1 = 1.0;
	if ((NodeType & NODE_BOUNDARY) != 0) w0 = 1.0/(3*nubuffer+0.5);""")
    second = ["200", "020", "002", "110", "101", "011"]
    third = ["210", "201", "120", "021", "102", "012", "111", "300"]
    third = [t for t in third if max(int(ch) for ch in t) <= 2]
    A("\treal_t " + ", ".join("c" + t for t in second + third) + ";")
    for t in second + third:
        A(f"\tc{t} = f{t}*irho;")
    A("""	vector_t u;
	u.x = u0.x + ForceX*0.5*irho;
	u.y = u0.y + ForceY*0.5*irho;
	u.z = u0.z + ForceZ*0.5*irho;""")
    if opt["SMAG"]:
        A(_SMAG)
    avg = "\tavgdxu2 = avgdxu2 + dxu*dxu;\n\tavgdyv2 = avgdyv2 + dyv*dyv;\n\tavgdzw2 = avgdzw2 + dzw*dzw;\n" if opt["AVG"] else ""
    A(_RELAX_2ND % dict(avg=avg))
    A("\treal_t om3 = 1-Omega;")
    for t in third:
        A(f"\tc{t} = om3*c{t};")
    # closed-form closure: central moments of order 4..6 with vanishing cumulants (normalised by rho)
    A("""	real_t c211 = c200*c011 + 2.0*c110*c101;
	real_t c121 = c020*c101 + 2.0*c110*c011;
	real_t c112 = c002*c110 + 2.0*c101*c011;
	real_t c220 = c200*c020 + 2.0*c110*c110;
	real_t c202 = c200*c002 + 2.0*c101*c101;
	real_t c022 = c020*c002 + 2.0*c011*c011;
	real_t c122 = c002*c120 + c020*c102 + 4.0*c011*c111 + 2.0*(c101*c021 + c110*c012);
	real_t c212 = c002*c210 + c200*c012 + 4.0*c101*c111 + 2.0*(c011*c201 + c110*c102);
	real_t c221 = c020*c201 + c200*c021 + 4.0*c110*c111 + 2.0*(c011*c210 + c101*c120);
	real_t c222 = 4.0*c111*c111 + c200*c022 + c020*c202 + c002*c220
		+ 4.0*(c011*c211 + c101*c121 + c110*c112)
		+ 2.0*(c120*c102 + c210*c012 + c201*c021)
		- 16.0*c110*c101*c011 - 4.0*(c101*c101*c020 + c011*c011*c200 + c110*c110*c002) - 2.0*c200*c020*c002;""")
    allm = [f"{a}{b}{c}" for c in range(3) for b in range(3) for a in range(3)]
    for t in allm:
        o = sum(int(ch) for ch in t)
        if o == 0:
            continue
        if o == 1:
            A(f"\tf{t} = 0.0;")
        else:
            A(f"\tf{t} = c{t}*rho;")
    A("""	vector_t un;      // the body force shifts the first cumulant: the backward sweep is taken about u0 + F
	un.x = u0.x + ForceX;
	un.y = u0.y + ForceY;
	un.z = u0.z + ForceZ;""")
    A(central_moment_backward("f", ["un.x", "un.y", "un.z"]))
    if opt["AVG"]:
        A(_AVG_TAIL)
    A("}")
