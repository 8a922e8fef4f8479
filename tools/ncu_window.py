#!/usr/bin/env python
"""Run a fixed list of kernels inside one cudaProfilerStart/Stop window, for ONE ncu invocation per GPU-box visit:

    python tools/ncu_window.py                       # plain run (must exit 0 before the ncu run)
    ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/prof_window \
        python tools/ncu_window.py

Window contents (512x256x256 fp32 unless noted; every launch is warmed up before the window opens):
  d3q27_cumulant   ring kernel with Globals (k_stage_persist<0,1,1> + k_fold_globals<1> + k_publish_globals),
                   classic kernel with Globals (k_stage<0,1>), classic without Globals (k_stage<0,0>),
                   k_quantity (U, P), k_color
  ref_d3q27_cumulant, ref_d3q19   classic kernel without Globals (the lowered reference models)
"""
import ctypes
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tclb_b200.runtime import Domain, Model  # noqa: E402

SHAPE = tuple(int(v) for v in os.environ.get("NCU_SHAPE", "512,256,256").split(","))


def cudart():
    for name in ("libcudart.so", "libcudart.so.12"):
        try:
            return ctypes.CDLL(name)
        except OSError:
            pass
    import torch  # the wheel bundles the runtime
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    for f in sorted(os.listdir(libdir)):
        if f.startswith("libcudart"):
            return ctypes.CDLL(os.path.join(libdir, f))
    raise RuntimeError("no libcudart")


def channel(name, shape):
    m = Model(name, "float")
    nx, ny, nz = shape
    d = Domain(m, shape, "cuda", nranks=1)
    fl = np.full((nz, ny, nx), m.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall")
    fl[:, -1, :] = m.flag("Wall")
    d.set_flags(fl)
    for k, v in (("nu", 0.01), ("ForceX", 1e-6)):
        if k in m.settings or k in m.zonesettings:
            d.set_setting(k, v)
    d.init()
    return m, d


def main():
    rt = cudart()
    todo = []
    os.environ["TCLB_KERNEL"] = "ring"
    m, d = channel("d3q27_cumulant", SHAPE)
    d.iterate(3, globals_every=True)
    d.get_quantity("U"); d.get_quantity("P"); d.color(SHAPE[2] // 2)
    d.sync()

    def native():
        os.environ["TCLB_KERNEL"] = "ring"
        d.iterate(1, globals_every=True)
        os.environ["TCLB_KERNEL"] = "classic"
        d.iterate(1, globals_every=True)
        d.iterate(1)
        d.get_quantity("U"); d.get_quantity("P"); d.color(SHAPE[2] // 2)
        d.sync()

    os.environ["TCLB_KERNEL"] = "classic"
    d.iterate(2, globals_every=True); d.iterate(2); d.sync()
    todo.append(native)
    keep = [d]
    for name in ("ref_d3q27_cumulant", "ref_d3q19"):
        try:
            m2, d2 = channel(name, SHAPE)
        except Exception as e:  # a library that is not built is reported, the others are still captured
            print("skip", name, e)
            continue
        d2.iterate(3); d2.sync()
        keep.append(d2)
        todo.append(lambda dd=d2: (dd.iterate(1), dd.sync()))
    # the four stages of one d3q27_pf_velocity iteration (lowered from the reference's files)
    try:
        mp = Model("ref_d3q27_pf_velocity", "float")
        dp = Domain(mp, (256, 128, 128), "cuda", nranks=1)
        fl = np.zeros((128, 128, 256), dtype=np.uint32) | mp.flag("MRT")
        fl[:, 0, :] = mp.flag("Wall"); fl[:, -1, :] = mp.flag("Wall")
        dp.set_flags(fl)
        for k, v in (("Density_h", 1.0), ("Density_l", 0.001), ("PhaseField_init", 1.0), ("sigma", 1e-3), ("M", 0.05), ("IntWidth", 4.0),
                     ("Viscosity_l", 0.16), ("Viscosity_h", 0.16), ("Radius", 30.0), ("CenterX", 128.0), ("CenterY", 64.0), ("CenterZ", 64.0), ("BubbleType", -1.0)):
            if k in mp.settings or k in mp.zonesettings:
                dp.set_setting(k, v)
        dp.init(); dp.iterate(3); dp.sync()
        keep.append(dp)
        todo.append(lambda: (dp.iterate(1), dp.sync()))
    except Exception as e:
        print("skip ref_d3q27_pf_velocity", e)
    rt.cudaProfilerStart()
    for f in todo:
        f()
    rt.cudaProfilerStop()
    for dd in keep:
        dd.close()
    print("window done")


if __name__ == "__main__":
    main()
