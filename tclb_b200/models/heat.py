"""Heat-transfer models.

* ``d3q19_heat`` - d3q19 flow lattice + d3q7 advected temperature lattice, heater node type
  (reference models/heat/d3q19_heat: Dynamics.R:1-40, Dynamics.c.Rt 177 lines).
"""
from __future__ import annotations

import numpy as np

from ..gen.lbmlib import d3q19_velocities, full_bounce_back, mrt_eq
from ..gen.modeldef import ModelDef
from ..gen.poly import PV, Poly, PVec, c_assign, to_c
from . import ModelSpec

D3Q7 = np.array([[0, 0, 0], [1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], dtype=int)


def build_d3q19_heat(options=None) -> ModelDef:
    m = ModelDef("d3q19_heat", dict(options or {}))
    U = d3q19_velocities()
    fn = [f"f{i}" for i in range(19)]
    gn = [f"g{i}" for i in range(7)]
    for i, (n, u) in enumerate(zip(fn, U)):
        m.AddDensity(name=n, dx=int(u[0]), dy=int(u[1]), dz=int(u[2]), comment=f"flow LB density F {i}", group="f")
    for i, (n, u) in enumerate(zip(gn, D3Q7)):
        m.AddDensity(name=n, dx=int(u[0]), dy=int(u[1]), dz=int(u[2]), comment=f"heat LB density G {i}", group="g")
    m.AddQuantity(name="Rho")
    m.AddQuantity(name="T")
    m.AddQuantity(name="U", vector=True)
    m.AddSetting(name="nu", default=0.16666666, comment="viscosity")
    m.AddSetting(name="Velocity", default="0m/s", comment="inlet velocity", zonal=True)
    m.AddSetting(name="Pressure", default="0Pa", comment="inlet pressure", zonal=True)
    m.AddSetting(name="Temperature", default=1, comment="inlet density", zonal=True)
    m.AddSetting(name="FluidAlpha", default=1, comment="inlet density")
    m.AddNodeType(name="Heater", group="ADDITIONALS")
    for n in ["EPressure", "EVelocity", "Solid", "Wall", "WPressure", "WVelocity"]:
        m.AddNodeType(name=n, group="BOUNDARY")
    m.AddNodeType(name="BGK", group="COLLISION")
    m.AddNodeType(name="MRT", group="COLLISION")
    f, g = PV(fn), PV(gn)
    rho, rhoT = Poly.sym("rho"), Poly.sym("rhoT")
    u = PV(["ux", "uy", "uz"])
    feq = mrt_eq(U, rho, PVec([x * rho for x in u]))
    geq = mrt_eq(D3Q7, rhoT, PVec([x * rhoT for x in u]), order=1, sigma2=1.0 / 4.0)
    jx, jy, jz = (f @ U[:, k].astype(float) for k in range(3))
    c = []
    A = c.append
    A(f"CudaDeviceFunction real_t getRho() {{\n\treturn {to_c(f.sum())};\n}}")
    A(f"CudaDeviceFunction real_t getT() {{\n\treturn ({to_c(g.sum())})/({to_c(f.sum())});\n}}")
    A("CudaDeviceFunction vector_t getU() {\n\treal_t d = getRho();\n\tvector_t u;")
    A(f"\tu.x = ({to_c(jx)})/d;\n\tu.y = ({to_c(jy)})/d;\n\tu.z = ({to_c(jz)})/d;\n\treturn u;\n}}")
    A("""CudaDeviceFunction float2 Color() {
	float2 ret;
	vector_t u = getU();
	ret.x = sqrt(u.x*u.x + u.y*u.y + u.z*u.z);
	ret.y = (NodeType == NODE_Solid) ? 0 : 1;
	return ret;
}""")
    A("CudaDeviceFunction void BounceBack() {\n" + full_bounce_back(m.density_dicts()) + "}")
    A("""CudaDeviceFunction void Run() {
	switch (NodeType & NODE_BOUNDARY) {
	case NODE_Wall:
	case NODE_Solid: BounceBack(); break;
	}
	switch (NodeType & NODE_COLLISION) {
	case NODE_MRT: CollisionMRT(); break;
	}
}""")
    A("CudaDeviceFunction void SetEquilibrum(real_t rho, real_t ux, real_t uy, real_t uz, real_t rhoT) {")
    A(c_assign(f, feq.feq))
    A(c_assign(g, geq.feq))
    A("}")
    A("CudaDeviceFunction void Init() {\n\tSetEquilibrum(1.0, Velocity, 0., 0., Temperature*1.0);\n}")
    A("CudaDeviceFunction void CollisionMRT() {\n\treal_t M[19];\n\treal_t rho, ux, uy, uz, rhoT, Q;")
    R = PV("M[", range(19), "]")
    A(f"\trho = {to_c(f.sum())};")
    A(f"\tux = ({to_c(jx)})/rho;\n\tuy = ({to_c(jy)})/rho;\n\tuz = ({to_c(jz)})/rho;")
    A(f"\trhoT = {to_c(g.sum())};")
    A("\tQ = 0;\n\tif ((NodeType & NODE_ADDITIONALS) == NODE_Heater) Q = Temperature*rho - rhoT;")
    A("\treal_t omega = 1.0-1.0/(3*nu + 0.5);\n\treal_t omegaT = 1.0-1.0/(3*FluidAlpha + 0.5);")
    A(c_assign(R, ((f - feq.feq) @ feq.mat) * Poly.sym("omega")))
    A(c_assign(R, R + feq.Req))
    nf = (feq.mat ** 2).sum(axis=0)
    A(c_assign(R, R * PVec([Poly.const(1.0 / v) for v in nf])))
    A(c_assign(f, R @ feq.mat.T))
    RT = R[:7]
    A(c_assign(RT, ((g - geq.feq) @ geq.mat) * Poly.sym("omegaT")))
    A("\trhoT += Q;")
    A(c_assign(RT, RT + geq.Req))
    ng = (geq.mat ** 2).sum(axis=0)
    A(c_assign(RT, RT * PVec([Poly.const(1.0 / v) for v in ng])))
    A(c_assign(g, RT @ geq.mat.T))
    A("}")
    m.code = "\n".join(c)
    return m


MODELS = {
    "d3q19_heat": ModelSpec("d3q19_heat", build_d3q19_heat, reference="models/heat/d3q19_heat"),
}
