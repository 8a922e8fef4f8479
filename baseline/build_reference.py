#!/usr/bin/env python
"""Builds the UNMODIFIED reference (nkkko/TCLB) for one model into baseline/_ref/CLB/<model>/main.

The reference generates its sources with R (`tools/RT` + the R packages rtemplate / polyAlgebra / gvector) and
builds them with autoconf + MPI; none of that exists in this image.  What this script does instead:

  1. expands every `src/*.Rt` (and the model's `Dynamics.c.Rt`) with this repository's R-subset interpreter and
     template engine (`tclb_b200.gen.rinterp/rtemplate`): the reference's own `src/conf.R` is *executed*, nothing is
     restated.  The expanded text is written to baseline/_ref/CLB/<model>/ next to verbatim copies of the
     non-templated sources - exactly the file set `make <model>` produces (src/makefile.main.Rt:232-296);
  2. writes the `config.h` / `config.mk` values that `./configure --enable-cuda --with-cuda-arch=sm_100` (float)
     gives (src/configure.ac);
  3. compiles with nvcc/g++ using the reference's own flags (src/config.mk.in, src/makefile.Rt) against a
     single-node stand-in for MPI (baseline/mpi_stub: the ~40 entry points the reference calls, implemented over
     POSIX shared memory so `-np N` ranks on one box work).

Only the generator and the MPI library are substituted; every kernel, the host engine, the handlers and the XML
front-end that run are the reference's own code.  None of tclb_b200's kernels, engine or models is involved.
"""
from __future__ import annotations

import argparse
import os
import re
import shutil
import subprocess
import sys
import time
from collections import OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
REF = Path(os.environ.get("TCLB_REFERENCE", "/root/reference"))
OUT = ROOT / "baseline" / "_ref"
CUDA_HOME = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))

MODEL_DIRS = {"d3q27_cumulant": "flow/d3q27_cumulant", "d2q9": "flow/d2q9", "d3q19": "flow/d3q19",
              "d3q27_pf_velocity": "multiphase/d3q27_pf_velocity", "d3q19_heat": "heat/d3q19_heat",
              "d2q9_poison_boltzmann": "electrokinetic/d2q9_poison_boltzmann"}


def default_options(model: str, on=()) -> "OrderedDict":
    """Options list of CLB/<model>/options.R: every option named in the model's conf.mk, switched on if listed in `on`
    (src/models.R:28-66, src/makefile.main.Rt:143-163)"""
    from tclb_b200.gen.rfrontend import all_option_names, parse_conf_mk
    conf = parse_conf_mk(str(REF / "models" / MODEL_DIRS[model] / "conf.mk"))
    opts = OrderedDict((o, (o in on)) for o in all_option_names(conf.get("OPT", "")))
    if "autosym" in opts:
        opts["autosym"] = 1 if "autosym" in on else 0
    return opts


def expand_one(model: str, template: str, on=()) -> str:
    """text of one expanded reference template (e.g. "LatticeContainer.inc.cpp.Rt") for `model`"""
    model_dir = REF / "models" / MODEL_DIRS[model]
    src = REF / "src"
    return render(model, model_dir, default_options(model, on), src / template, [REF / "tools", src, model_dir])


def make_interp(model: str, model_dir: Path, options: dict, adjoint: int = 0):
    from tclb_b200.gen import rlib
    from tclb_b200.gen.rinterp import Interp
    from tclb_b200.gen.rlang import RList, vec
    src = REF / "src"
    I = Interp(include_dirs=[str(model_dir), str(src), str(src / "lib"), str(REF / "tools")])
    I.plain_c = True          # C() prints assignments as they are written (no rewriting of this repository's generator)
    rlib.install(I)
    env = I.globals
    # CLB/<model>/options.R = CLB/config.R (X_MOD, CPU_LAYOUT from configure) + MODEL/ADJOINT/Options (makefile.main.Rt:143-163)
    env.set("X_MOD", vec(0))
    env.set("CPU_LAYOUT", vec(False))
    env.set("Options", RList.make([vec(v) for v in options.values()], list(options.keys())))
    env.set("MODEL", vec(model))
    if adjoint:
        env.set("ADJOINT", vec(1))
    env.set("include.dir", vec(str(model_dir)))
    return I


def render(model: str, model_dir: Path, options: dict, template: Path, search) -> str:
    from tclb_b200.gen import poly, rtemplate
    poly._tmp_counter[0] = 0
    I = make_interp(model, model_dir, options)
    try:
        return rtemplate.render(I, I.globals, str(template), [str(s) for s in search])
    except Exception as e:
        raise RuntimeError(f"{template}: {type(e).__name__}: {e}  [R call stack, innermost first: {getattr(e, 'rstack', None)}]") from e


CONFIG_H = """#ifndef CONFIG_H
  #define CONFIG_H
/* what ./configure --enable-cuda --with-cuda-arch=sm_100 (float, defaults otherwise) defines: src/configure.ac */
#define X_BLOCK 32
#define WARPSIZE 32
#define SOLID_CONTAINER 3
#define GRID3D 1
#define CUDA_CC "100"
#define CROSS_MPI_WAITANY 1
#define DEBUG_LEVEL 3
#define PACKAGE_NAME "CLB"
#define PACKAGE_STRING "CLB 6.0"
#define PACKAGE_TARNAME "clb"
#define PACKAGE_URL ""
#define PACKAGE_VERSION "6.0"
#define USE_STEADY_CLOCK 1
#define HAS_CXXABI_H 1
#endif
"""


def sp_const(text: str) -> str:
    """%_sp.c : %.c rule of src/makefile.main.Rt:293-296"""
    return re.sub(r"([^a-zA-Z0-9][0-9]+\.[0-9]*([eE][-+]?[0-9]+)?)[fL]+", r"\1f", text)


def expand(model: str, options: dict, verbose=True) -> Path:
    model_dir = REF / "models" / MODEL_DIRS[model]
    dst = OUT / "CLB" / model
    if dst.exists():
        shutil.rmtree(dst)
    (dst / "Handlers").mkdir(parents=True)
    src = REF / "src"
    search = [REF / "tools", src, model_dir]
    skip = {"makefile.main.Rt", "Makefile.in.Rt", "config.R.in", "config.h.in", "config.mk.in", "config.main.mk.in", "configure.ac",
            "Model.md.Rt", "Models.md.Rt", "SUMMARY.Rt", "schema.xsd.Rt", "catalog.xml.Rt", "Fast-Transforms.md.Rt", "solver.vcproj",
            "ADpost.sed.Rt", "ADpre.h.Rt", "ADpre_b.h.Rt", "ADset.sh.Rt", "Catalyst.cpp.Rt", "dictionary.h.Rt", "makefile.Rt", "conf.R", "main.R", "models.R", "README.md"}
    files = [p for p in sorted(src.iterdir()) if p.is_file() and p.name not in skip] + sorted((src / "Handlers").iterdir())
    t0 = time.time()
    for p in files:
        rel = p.relative_to(src)
        if p.suffix == ".Rt":
            out = dst / str(rel)[:-3]
            text = render(model, model_dir, options, p, search)
            out.write_text(text)
            if verbose:
                print(f"  RT   {rel} -> {out.relative_to(OUT)} ({len(text)} bytes, {time.time() - t0:.1f}s)", flush=True)
        else:
            shutil.copy(p, dst / rel)
    for p in sorted(model_dir.iterdir()):
        if p.suffix == ".Rt":
            text = render(model, model_dir, options, p, search)
            (dst / p.name[:-3]).write_text(text)
            if verbose:
                print(f"  RT   (model) {p.name} ({len(text)} bytes)", flush=True)
        elif p.is_file():
            shutil.copy(p, dst / p.name)
    (dst / "Dynamics_sp.c").write_text(sp_const((dst / "Dynamics.c").read_text()))
    (dst.parent / "config.h").write_text(CONFIG_H)
    return dst


OBJ = ["vtkOutput", "cuda", "Global", "Lattice", "vtkLattice", "cross", "pugixml", "Geometry", "def", "unit", "Solver", "SyntheticTurbulence",
       "Sampler", "ZoneSettings", "RemoteForceInterface", "hdf5Lattice", "xpath_modification", "GetThreads", "Lists", "main"]   # src/config.mk.in OBJ + main


def compile_model(dst: Path, jobs: int = 0, verbose=True) -> Path:
    """nvcc/g++ with the flags of src/config.mk.in + src/makefile.Rt for a `--enable-cuda --with-cuda-arch=sm_100` configuration --disable-cpp11"""
    import concurrent.futures as cf
    stub = ROOT / "baseline" / "mpi_stub"
    # --disable-cpp11: the CUDA 12.9 headers (libcu++ behind cooperative_groups.h) no longer compile as C++11, so no
    # -std flag is passed and both compilers use their default (C++17)
    cpp_opt = ["-I", str(stub), "-I", str(CUDA_HOME / "include"), "-O3", "-Wno-write-strings"]                      # CPPFLAGS
    nv_opt = ["-D_FORCE_INLINES", "-arch=compute_100", "--disable-warnings", "--compiler-options=" + " ".join(cpp_opt[4:]),
              "-I", str(stub),
              # src/cross.hpp:14-16,88-91 includes <cooperative_groups.h> only when __CUDA_ARCH__ >= 700, i.e. only in the
              # device pass; with CUDA 12.9 that header defines device variables in namespace cuda::std which the host
              # stub then cannot name.  Pre-including the headers (a compiler flag, no source change) declares them in
              # both passes.
              "--pre-include", "cooperative_groups.h", "--pre-include", "cooperative_groups/reduce.h"]            # NVFLAGS
    ld_opt = ["-L", str(CUDA_HOME / "lib64"), "-lcudart", "-lpthread", "-ldl", "-lrt"]
    handlers = sorted(p.stem for p in (dst / "Handlers").glob("*.cpp"))
    units = [(n, dst / f"{n}.cu" if (dst / f"{n}.cu").exists() else dst / f"{n}.cpp") for n in OBJ] + [("Handlers/" + h, dst / "Handlers" / f"{h}.cpp") for h in handlers]

    # nvcc 12.9 cannot compile a translation unit whose FILE NAME is cuda.cu (its host stub then refers to `::cuda`);
    # the file is compiled through a one-line wrapper instead, its text is untouched
    (dst / "tu_cuda.cu").write_text('#include "cuda.cu"\n')
    units = [(n, dst / "tu_cuda.cu") if n == "cuda" else (n, p_) for n, p_ in units]

    def one(u):
        name, src = u
        obj = dst / (name + ".o")
        if src.suffix == ".cu":
            cmd = [str(CUDA_HOME / "bin" / "nvcc"), "-c", str(src), "-o", str(obj)] + nv_opt
        else:
            cmd = ["g++", "-c", str(src), "-o", str(obj)] + cpp_opt
        t0 = time.time()
        r = subprocess.run(cmd, cwd=dst, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"compiling {src.name} failed:\n{' '.join(cmd)}\n{r.stdout[-4000:]}")
        if verbose:
            print(f"  {'NVCC' if src.suffix == '.cu' else 'C++ '} {name} ({time.time() - t0:.0f}s)", flush=True)
        return obj

    with cf.ThreadPoolExecutor(max_workers=jobs or os.cpu_count() or 4) as ex:
        objs = list(ex.map(one, units))
    stub_o = dst / "mpi_stub.o"
    subprocess.run(["g++", "-c", "-O2", "-std=c++11", str(stub / "mpi_stub.cpp"), "-I", str(stub), "-o", str(stub_o)], check=True)
    main = dst / "main"
    r = subprocess.run(["g++"] + [str(o) for o in objs] + [str(stub_o), "-o", str(main)] + ld_opt, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout[-4000:])
    return main


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("model", nargs="?", default="d3q27_cumulant")
    ap.add_argument("--expand-only", action="store_true")
    a = ap.parse_args()
    d = expand(a.model, default_options(a.model))
    print("expanded into", d)
    if not a.expand_only:
        print("built", compile_model(d))
