"""Electrokinetic models.

* ``d2q9_poison_boltzmann`` - Poisson-Boltzmann potential solved by a d2q9 diffusion lattice with a
  sinh source, three-stage iteration (collision+stream, potential field, sub-iteration counter)
  (reference models/electrokinetic/d2q9_poison_boltzmann: Dynamics.R 73 lines, Dynamics.c.Rt 116 lines).
"""
from __future__ import annotations

import numpy as np

from ..gen.lbmlib import d2q9_velocities
from ..gen.modeldef import ModelDef
from ..gen.poly import PV, Poly, PVec, c_assign, to_c
from . import ModelSpec


def build_poison_boltzmann(options=None) -> ModelDef:
    m = ModelDef("d2q9_poison_boltzmann", dict(options or {}))
    m.SetOptions(permissive_access=True)
    U = d2q9_velocities()
    for i, u in enumerate(U):
        m.AddDensity(name=f"g[{i}]", dx=int(u[0]), dy=int(u[1]), group="g")
    m.AddDensity(name="subiter")
    m.AddField(name="psi", stencil2d=1)
    m.AddQuantity(name="Psi")
    m.AddQuantity(name="Subiter")
    m.AddStage("BaseIteration", "Run", save_fields=[f.name for f in m.fields if f.group == "g"],
               load_densities=[d.name for d in m.densities if d.group == "g"])
    m.AddStage("CalcPsi", save_fields=["psi"], load_densities=[d.name for d in m.densities if d.group == "g"])
    m.AddStage("CalcSubiter", save_fields=["subiter"], load_densities=["subiter"])
    m.AddAction("Iteration", ["BaseIteration", "CalcPsi", "CalcSubiter"])
    m.AddQuantity(name="rho_e", unit="kg/m3")
    for n in ["tau_psi", "n_inf", "z", "el", "kb", "T", "epsilon", "dt"]:
        m.AddSetting(name=n, comment=n if n == "tau_psi" else "")
    m.AddSetting(name="psi_bc", default=1, comment="psi at  boundary - zeta", zonal=True)
    m.AddSetting(name="psi0", default=1, comment="initial psi - zeta", zonal=True)
    m.AddNodeType(name="Solid", group="BOUNDARY")
    m.AddNodeType(name="Wall", group="BOUNDARY")
    m.AddNodeType(name="Collision", group="COLLISION")
    g = PV("g[", range(9), "]")
    wp = np.array([1. / 9. - 1.] + [1. / 9.] * 8)
    wps = np.array([0.] + [1. / 8.] * 8)
    wp0 = 1. / 9.
    psi, rho_e, tau, dt, eps = (Poly.sym(s) for s in ["psi", "rho_e", "tau_psi", "dt", "epsilon"])
    RD = (Poly.const(0.5) - tau) * dt * rho_e * eps ** -1 * (-2. / 3.)
    c = []
    A = c.append
    A(f"CudaDeviceFunction real_t getPsi() {{\n\tconst real_t wp0 = {wp0!r};\n\treturn ( {to_c(PVec(g.v[1:]).sum())} ) / (1. - wp0);\n}}")
    A("CudaDeviceFunction real_t getSubiter() { return subiter; }")
    A("CudaDeviceFunction void CalcPsi() { psi = getPsi(); }")
    A("""CudaDeviceFunction real_t getrho_e() {
	const real_t psi = getPsi();
	return -2. * n_inf * z * el * sinh( z * el / kb / T * psi);
}""")
    A("CudaDeviceFunction void BounceBack() {")
    A(c_assign(g, PVec([Poly.sym("psi_bc") * float(w) for w in wp])))
    A("}")
    A("CudaDeviceFunction float2 Color() {\n\tfloat2 ret;\n\tret.x = 0;\n\tret.y = 0;\n\treturn ret;\n}")
    A("""CudaDeviceFunction void Run() {
	switch (NodeType & NODE_BOUNDARY) {
	case NODE_Solid:
	case NODE_Wall: BounceBack(); break;
	}
	if (NodeType & NODE_COLLISION) CollisionBGK();
}""")
    A("CudaDeviceFunction void Init() {")
    A(c_assign(g, PVec([Poly.sym("psi0") * float(w) for w in wp])))
    A("\tsubiter = 0;\n\tpsi = psi0;\n}")
    A("CudaDeviceFunction void CollisionBGK() {\n\treal_t psi, rho_e;\n\trho_e = getrho_e();\n\tpsi = getPsi();")
    newg = PVec([g[i] - tau ** -1 * (g[i] - psi * float(wp[i])) + dt * RD * float(wps[i]) for i in range(9)])
    A(c_assign(g, newg))
    A("}")
    A("CudaDeviceFunction void CalcSubiter() { subiter = subiter + 1; }")
    m.code = "\n".join(c)
    return m


MODELS = {
    "d2q9_poison_boltzmann": ModelSpec("d2q9_poison_boltzmann", build_poison_boltzmann,
                                       reference="models/electrokinetic/d2q9_poison_boltzmann"),
}
