"""Finite-difference PDE models on the lattice engine (no populations, only neighbour-read fields):
diffusion2D, wave2D, advection_diffusion2D[_fields].  Same surface as the reference's
``models/PDE/{diffusion2D,wave2D,advection_diffusion2D}`` (each ~50 lines of Dynamics.R + Dynamics.c)."""
from __future__ import annotations

from ..gen.modeldef import ModelDef
from . import ModelSpec

_COLOR = """CudaDeviceFunction float2 Color() {
	float2 ret;
	ret.x = get%s();
	ret.y = 1;
	return ret;
}
"""


def build_diffusion2d(options=None) -> ModelDef:
    m = ModelDef("diffusion2D", dict(options or {}))
    m.AddField(name="phi", dx=[-1, 1], dy=[-1, 1])
    m.AddQuantity(name="Phi")
    m.AddSetting(name="diff_coeff")
    m.AddSetting(name="Value", zonal=True)
    m.AddNodeType(name="Dirichlet", group="BOUNDARY")
    m.code = _COLOR % "Phi" + """
CudaDeviceFunction real_t getPhi() { return phi(0,0); }
CudaDeviceFunction void Init() { phi = Value; }
CudaDeviceFunction void Run() {
	real_t lap_phi = phi(-1,0) + phi(1,0) + phi(0,-1) + phi(0,1) - 4*phi(0,0);
	phi = phi(0,0) + diff_coeff * diff_coeff * lap_phi;
	if ((NodeType & NODE_BOUNDARY) == NODE_Dirichlet) phi = Value;
}
"""
    return m


def build_wave2d(options=None) -> ModelDef:
    m = ModelDef("wave2D", dict(options or {}))
    m.AddField(name="u", dx=[-1, 1], dy=[-1, 1])
    m.AddField(name="v", dx=[-1, 1], dy=[-1, 1])
    m.AddQuantity(name="U")
    m.AddSetting(name="Speed")
    m.AddSetting(name="Viscosity")
    m.AddSetting(name="Value", zonal=True)
    m.AddNodeType(name="Dirichlet", group="BOUNDARY")
    m.code = _COLOR % "U" + """
CudaDeviceFunction real_t getU() { return u(0,0); }
CudaDeviceFunction void Init() { u = Value; v = 0; }
CudaDeviceFunction void Run() {
	real_t lap_u = u(-1,0) + u(1,0) + u(0,-1) + u(0,1) - 4*u(0,0);
	real_t lap_v = v(-1,0) + v(1,0) + v(0,-1) + v(0,1) - 4*v(0,0);
	real_t acc = Speed * Speed * lap_u + Viscosity * lap_v;
	v = v(0,0) + acc;
	u = u(0,0) + v;
	if ((NodeType & NODE_BOUNDARY) == NODE_Dirichlet) { u = Value; v = 0; }
}
"""
    return m


def build_advection_diffusion2d(options=None) -> ModelDef:
    opt = dict(fields=False)
    opt.update(options or {})
    m = ModelDef("advection_diffusion2D", opt)
    m.AddField(name="phi", dx=[-1, 1], dy=[-1, 1])
    m.AddQuantity(name="Phi")
    m.AddSetting(name="diff_coeff")
    m.AddSetting(name="Value", zonal=True)
    if opt["fields"]:
        m.AddDensity(name="ux", group="u", comment="free stream velocity", parameter=True)
        m.AddDensity(name="uy", group="u", comment="free stream velocity", parameter=True)
        m.AddDensity(name="phi0", group="init", comment="initial phi", parameter=True)
        m.AddStage(name="InitFromFieldsStage", main="InitFromFields", load_densities=True, save_fields=True)
        m.AddAction(name="InitFromFields", stages=["InitFromFieldsStage"])
    else:
        m.AddSetting(name="ux", comment="free stream velocity")
        m.AddSetting(name="uy", comment="free stream velocity")
    m.AddNodeType(name="Dirichlet", group="BOUNDARY")
    init_extra = "\tux = 0;\n\tuy = 0;\n\tphi0 = 0;\n" if opt["fields"] else ""
    m.code = _COLOR % "Phi" + """
CudaDeviceFunction real_t getPhi() { return phi(0,0); }
CudaDeviceFunction void Init() {
	phi = Value;
%s}
%s
CudaDeviceFunction void Run() {
	real_t lap_phi = phi(-1,0) + phi(1,0) + phi(0,-1) + phi(0,1) - 4*phi(0,0);
	real_t gx = (phi(1,0)-phi(-1,0))/2;
	real_t gy = (phi(0,1)-phi(0,-1))/2;
	phi = phi(0,0) + diff_coeff * diff_coeff * lap_phi - ux*gx - uy*gy;
	if ((NodeType & NODE_BOUNDARY) == NODE_Dirichlet) phi = Value;
}
""" % (init_extra, "CudaDeviceFunction void InitFromFields() { phi = phi0; }" if opt["fields"] else "")
    return m


MODELS = {
    "diffusion2D": ModelSpec("diffusion2D", build_diffusion2d, reference="models/PDE/diffusion2D"),
    "wave2D": ModelSpec("wave2D", build_wave2d, reference="models/PDE/wave2D"),
    "advection_diffusion2D": ModelSpec("advection_diffusion2D", build_advection_diffusion2d, ["fields"], [dict(), dict(fields=True)],
                                       reference="models/PDE/advection_diffusion2D"),
}
