import numpy as np


def channel_flags(model, shape, obstacle=True):
    nx, ny, nz = shape
    fl = np.full((nz, ny, nx), model.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = model.flag("Wall")
    fl[:, -1, :] = model.flag("Wall")
    if obstacle:
        z0, z1 = (0, 1) if nz == 1 else (nz // 3, nz // 3 + max(2, nz // 5))
        fl[z0:z1, ny // 3:ny // 3 + max(2, ny // 5), nx // 4:nx // 4 + 4] = model.flag("Wall")
    return fl


def setup_case(domain, model, shape, force=1e-5, visc=0.05):
    domain.set_flags(channel_flags(model, shape))
    if "nu" in model.settings:
        domain.set_setting("nu", visc)
        domain.set_setting("ForceX", force)
    else:
        domain.set_setting("Viscosity", visc)
        domain.set_setting("GravitationX", force)
    domain.init()


def run_case(model_name, precision, backend, shape, iters, nranks=1, direct=True, devices=None):
    from tclb_b200.runtime import Domain, Model
    m = Model(model_name, precision)
    d = Domain(m, shape, backend, nranks=nranks, devices=devices)
    if not direct:
        d.set_halo_mode(False)
    setup_case(d, m, shape)
    d.iterate(iters, globals_on_last=True)
    out = {q: d.get_quantity(q) for q in m.quantities}
    out["_globals"] = d.globals()
    out["_fields"] = {f: d.get_field(f) for f in m.fields[:3]}
    d.close()
    return out
