"""Build orchestration: generate model sources and compile every native piece for sm_100a.

Replaces the reference's configure + R-template + make pipeline (``makefile``,
``src/makefile.main.Rt:165-274``, ``src/configure.ac``): `python -m tclb_b200.build d2q9 d3q19`
produces ``tclb_b200/_lib/libtclb_core.so``, ``tclb_b200/_lib/libtclb_<model>_<float|double>.so``
and ``bin/tclb``.  Artefacts are built in-tree and keyed by a hash of sources+flags.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import json
import os
import threading
import shutil
import subprocess
import sys
from pathlib import Path
from typing import Iterable, List, Optional

ROOT = Path(__file__).resolve().parent.parent
LIBDIR = ROOT / "tclb_b200" / "_lib"
GENDIR = ROOT / "build" / "gen"
BINDIR = ROOT / "bin"
CUDA_HOME = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))
NVCC = str(CUDA_HOME / "bin" / "nvcc")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC,-fopenmp,-O2",
              "-Xptxas", "-v", "-diag-suppress", "177,550"] + os.environ.get("TCLB_NVCC_EXTRA", "").split()    # extra -D flags for A/B builds
CXX_FLAGS = ["-O2", "-g", "-std=c++17", "-fPIC", "-Wall", "-Wno-unused-function", "-fopenmp"]


def _run(cmd: List[str], log: Optional[Path] = None) -> str:
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.parent.mkdir(parents=True, exist_ok=True)
        log.write_text(" ".join(cmd) + "\n" + p.stdout)
    if p.returncode != 0:
        raise RuntimeError(f"command failed ({p.returncode}): {' '.join(cmd)}\n{p.stdout[-6000:]}")
    return p.stdout


def _hash(paths: Iterable[Path], extra: str = "") -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(str(p.relative_to(ROOT)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _stamp_ok(target: Path, digest: str) -> bool:
    st = target.with_suffix(target.suffix + ".stamp")
    return target.exists() and st.exists() and st.read_text() == digest


def _stamp(target: Path, digest: str) -> None:
    target.with_suffix(target.suffix + ".stamp").write_text(digest)


CORE_SOURCES = ["csrc/core", "csrc/host", "csrc/handlers", "csrc/io"]


def core_sources() -> List[Path]:
    out: List[Path] = []
    for d in CORE_SOURCES:
        out += sorted((ROOT / d).glob("*.cpp"))
    return out


def core_cu_sources() -> List[Path]:
    out: List[Path] = []
    for d in CORE_SOURCES + ["csrc/kernels"]:
        out += sorted((ROOT / d).glob("*.cu"))
    return out


def build_core(force: bool = False, verbose: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    target = LIBDIR / "libtclb_core.so"
    srcs = core_sources()
    cus = core_cu_sources()
    hdrs = [p for d in CORE_SOURCES + ["csrc/kernels", "include"] for p in (ROOT / d).glob("*.h")]
    digest = _hash(srcs + cus + hdrs, " ".join(CXX_FLAGS + NVCC_FLAGS))
    if not force and _stamp_ok(target, digest):
        return target
    objdir = ROOT / "build" / "core"
    objdir.mkdir(parents=True, exist_ok=True)
    incs = ["-I", str(ROOT / "include"), "-I", str(ROOT / "csrc"), "-I", str(ROOT / "csrc/core"), "-I", str(ROOT / "csrc/kernels"),
            "-I", str(CUDA_HOME / "include")]

    def compile_one(src: Path) -> Path:
        obj = objdir / (src.parent.name + "_" + src.stem + ".o")
        d = _hash([src] + hdrs, "obj")
        if not force and _stamp_ok(obj, d):
            return obj
        if src.suffix == ".cu":
            _run([NVCC] + ARCH_FLAGS + NVCC_FLAGS + incs + ["-c", str(src), "-o", str(obj)], objdir / (obj.name + ".log"))
        else:
            _run(["g++"] + CXX_FLAGS + incs + ["-c", str(src), "-o", str(obj)], objdir / (obj.name + ".log"))
        _stamp(obj, d)
        return obj

    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        objs = list(ex.map(compile_one, srcs + cus))
    _run(["g++", "-shared", "-o", str(target)] + [str(o) for o in objs] +
         ["-L", str(CUDA_HOME / "lib64"), "-lcudart_static", "-ldl", "-lrt", "-lpthread", "-lgomp", "-lz"])
    _stamp(target, digest)
    # command line driver
    main_src = ROOT / "csrc" / "main.cpp"
    if main_src.exists():
        BINDIR.mkdir(exist_ok=True)
        _run(["g++"] + CXX_FLAGS + incs + [str(main_src), "-o", str(BINDIR / "tclb"), "-L", str(LIBDIR), "-ltclb_core",
                                          "-Wl,-rpath,$ORIGIN/../tclb_b200/_lib", "-ldl"])
    integ_src = ROOT / "csrc" / "integrator_main.cpp"
    if integ_src.exists():
        _run(["g++"] + CXX_FLAGS + incs + [str(integ_src), "-o", str(BINDIR / "tclb-integrator"), "-L", str(LIBDIR), "-ltclb_core",
                                          "-Wl,-rpath,$ORIGIN/../tclb_b200/_lib", "-ldl"])
    cmp_src = ROOT / "csrc" / "compare.cpp"
    if cmp_src.exists():
        _run(["g++"] + CXX_FLAGS + incs + [str(cmp_src), "-o", str(BINDIR / "compare"), "-L", str(LIBDIR), "-ltclb_core",
                                          "-Wl,-rpath,$ORIGIN/../tclb_b200/_lib", "-ldl"])
    if verbose:
        print(f"built {target}")
    return target


_GEN_LOCK = threading.Lock()


def generate_model(variant: str) -> Path:
    from .gen.emit import emit_model_source
    from .models import build_model
    GENDIR.mkdir(parents=True, exist_ok=True)
    out = GENDIR / f"{variant}.cu"
    # one generation at a time: the symbolic layer numbers its temporaries with a per-model counter, and the
    # float and double builds of a model (separate threads in build_all) share this file
    with _GEN_LOCK:
        m = build_model(variant)
        src = emit_model_source(m, variant)
        if not out.exists() or out.read_text() != src:
            tmp = out.with_suffix(".cu.tmp")
            tmp.write_text(src)
            os.replace(tmp, out)
    return out


PRECISIONS = ("float", "double", "float_half", "float_halfshift", "double_float", "double_floatshift")


def precision_flags(precision: str) -> List[str]:
    """compile flags of a precision/storage combination: <real>[_<storage>] with storage in half, halfshift, float,
    floatshift (the reference's ./configure --with-storage=half|half-shift|float|float-shift next to --enable-double;
    plain float/double store what they compute)"""
    if precision not in PRECISIONS:
        raise ValueError(f"precision must be one of {PRECISIONS}")
    real, _, storage = precision.partition("_")
    flags = ["-DTCLB_DOUBLE"] if real == "double" else []
    if storage.startswith("half"):
        flags.append("-DTCLB_STORAGE_HALF")
    if storage.startswith("float"):
        flags.append("-DTCLB_STORAGE_FLOAT")
    if storage.endswith("shift"):
        flags.append("-DTCLB_STORAGE_SHIFT")
    return flags


def source_flags(src: Path) -> List[str]:
    """nvcc flags the generator asks for in the first lines of a generated source (`// nvcc-flags: ...`), e.g.
    -prec-div=false: fp32 divisions as reciprocal + multiply (2 ulp) instead of the IEEE sequence with its slow-path call;
    an IEEE division is ~20 issue slots, the reference's cumulant collision has 22 of them per node."""
    out: List[str] = []
    with open(src) as f:
        for _ in range(6):
            line = f.readline()
            if line.startswith("// nvcc-flags:"):
                out += line.split(":", 1)[1].split()
    return out


def build_model_lib(variant: str, precision: str = "float", force: bool = False, verbose: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    src = generate_model(variant)
    target = LIBDIR / f"libtclb_{variant}_{precision}.so"
    hdrs = [ROOT / "include" / "tclb_model_abi.h", ROOT / "csrc" / "kernels" / "tclb_node_rt.h"]
    digest = _hash([src] + hdrs, precision + " ".join(NVCC_FLAGS + ARCH_FLAGS))
    if not force and _stamp_ok(target, digest):
        return target
    cmd = [NVCC] + ARCH_FLAGS + NVCC_FLAGS + source_flags(src) + ["-shared", "-I", str(ROOT / "include"), "-I", str(ROOT / "csrc" / "kernels")]
    cmd += precision_flags(precision)
    cmd += ["-o", str(target), str(src), "-lgomp"]
    log = ROOT / "build" / "logs" / f"{variant}_{precision}.ptxas.log"
    _run(cmd, log)
    _stamp(target, digest)
    if verbose:
        print(f"built {target}")
    return target


_CUSTOM_MODELS: dict = {}      # library name -> how to load its ModelDef again (models built from a directory)


def build_adjoint_lib(variant: str, precision: str = "double", dirs: Optional[int] = None, force: bool = False, check_only: bool = False) -> Optional[Path]:
    """The adjoint companion of a model library: the same generated source compiled with -DTCLB_AD (node code in dual
    numbers, csrc/kernels/tclb_dual.h), exporting ``tclb_model_adjoint_stage``.  The runtime loads
    ``libtclb_<variant>_<precision>_ad.so`` next to the primal library when an adjoint is asked for.  `variant` is a
    native model name or the name of an already generated source (``ref_<model>`` after build_from_dir).
    ``check_only``: compile nothing; return the library if it matches the source the generator emits now, else None."""
    if precision not in ("float", "double"):
        raise ValueError("adjoint libraries use plain float/double storage")
    LIBDIR.mkdir(parents=True, exist_ok=True)
    GENDIR.mkdir(parents=True, exist_ok=True)
    from .gen.emit import emit_model_source
    src = GENDIR / f"{variant}_ad.cu"
    with _GEN_LOCK:
        if variant in _CUSTOM_MODELS:
            m = _CUSTOM_MODELS[variant]()
        else:
            from .models import build_model
            m = build_model(variant)
        text = emit_model_source(m, variant, adjoint=True)
        if dirs is None:
            # tangent directions per pass: every input at once for small velocity sets, 5 otherwise (measured on B200:
            # d2q9 13.5 GLUPS at 9 vs 10.2 at 4; d3q19 spills beyond 7) - profiles/README.md
            n_in = max((sum(1 for d in m.densities if d.load.get(s.name)) for s in m.stages), default=1)
            dirs = n_in if n_in <= 9 else 5
        if not src.exists() or src.read_text() != text:
            src.write_text(text)
    target = LIBDIR / f"libtclb_{variant}_{precision}_ad.so"
    hdrs = [ROOT / "include" / "tclb_model_abi.h", ROOT / "csrc" / "kernels" / "tclb_node_rt.h", ROOT / "csrc" / "kernels" / "tclb_dual.h"]
    extra = ["-DTCLB_AD", f"-DTCLB_AD_DIRS={int(dirs)}"]
    digest = _hash([src] + hdrs, precision + " ".join(NVCC_FLAGS + ARCH_FLAGS + extra))
    if not force and _stamp_ok(target, digest):
        return target
    if check_only:
        return None
    cmd = [NVCC] + ARCH_FLAGS + NVCC_FLAGS + extra + ["-shared", "-I", str(ROOT / "include"), "-I", str(ROOT / "csrc" / "kernels")]
    cmd += precision_flags(precision)
    cmd += ["-o", str(target), str(src), "-lgomp"]
    _run(cmd, ROOT / "build" / "logs" / f"{variant}_{precision}_ad.ptxas.log")
    _stamp(target, digest)
    return target


def build_custom(modeldef, libname: str, precision: str = "double", extra_flags=()) -> Path:
    """Compile a ModelDef object under an explicit library name (tests, experiments, user models)."""
    from .gen.emit import emit_model_source
    GENDIR.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    src = GENDIR / f"{libname}.cu"
    with _GEN_LOCK:          # float and double builds of one library share this file
        text = emit_model_source(modeldef, libname)
        if not src.exists() or src.read_text() != text:
            tmp = src.with_suffix(".cu.tmp")
            tmp.write_text(text)
            os.replace(tmp, src)
    target = LIBDIR / f"libtclb_{libname}_{precision}.so"
    hdrs = [ROOT / "include" / "tclb_model_abi.h", ROOT / "csrc" / "kernels" / "tclb_node_rt.h"]
    digest = _hash([src] + hdrs, precision + " ".join(NVCC_FLAGS + ARCH_FLAGS + list(extra_flags)))
    if _stamp_ok(target, digest):
        return target
    cmd = [NVCC] + ARCH_FLAGS + NVCC_FLAGS + source_flags(src) + list(extra_flags) + ["-shared", "-I", str(ROOT / "include"), "-I", str(ROOT / "csrc" / "kernels")]
    cmd += precision_flags(precision)
    cmd += ["-o", str(target), str(src), "-lgomp"]
    _run(cmd, ROOT / "build" / "logs" / f"{libname}_{precision}.ptxas.log")
    _stamp(target, digest)
    return target


DEFAULT_MODELS = ["d2q9", "d2q9_bc", "d3q19", "d3q27_cumulant"]


def all_variants() -> List[str]:
    from .models import base_models, variants
    out: List[str] = []
    for b in base_models():
        out += variants(b)
    return out


def build_from_dir(model_dir: str, name: Optional[str] = None, precision: str = "float", options: Optional[dict] = None,
                   libname: Optional[str] = None) -> Path:
    """Build a model library straight from an unmodified TCLB model directory (Dynamics.R + Dynamics.c[.Rt] + conf.mk):
    the R-subset interpreter and the template engine lower it, the generator emits the fused sm_100a kernels.
    The library is named ``ref_<model>[_<option>...]`` unless `libname` is given."""
    from .gen import rfrontend
    with _GEN_LOCK:
        m = rfrontend.load_model(model_dir, name=name, options=options)
    on = [k + (str(int(v)) if not isinstance(v, bool) and int(v) > 1 else "") for k, v in (options or {}).items() if v]
    lib = libname or "ref_" + "_".join([m.name] + on)
    _CUSTOM_MODELS[lib] = lambda: rfrontend.load_model(model_dir, name=name, options=options)
    return build_custom(m, lib, precision)


# models of a mounted TCLB tree that are built next to the native ones when the tree is present
REFERENCE_ROOT = Path(os.environ.get("TCLB_REFERENCE", "/root/reference"))
REFERENCE_MODELS = ["multiphase/d3q27_pf_velocity", "flow/d2q9", "flow/d3q19", "flow/d3q27_cumulant", "heat/d3q19_heat",
                    "electrokinetic/d2q9_poison_boltzmann", "electrokinetic/d2q9_npe_guo", "multiphase/d2q9_kuper",
                    "particles/d3q27_PSM:NEBB", "optimization/d2q9_adj"]            # "<directory>[:<option>,...]"
REFERENCE_ADJOINT = ["ref_d2q9_adj"]             # libraries that also get their adjoint companion (build_adjoint_lib)


def build_reference_models(precisions=("float", "double"), verbose=False, jobs=None) -> List[Path]:
    """Libraries lowered from the mounted reference tree's own model directories.  Front-end and source generation run one
    at a time (shared symbolic state, _GEN_LOCK); the nvcc compilations run side by side."""
    out: List[Path] = []
    if not (REFERENCE_ROOT / "models").is_dir():
        return out
    tasks = []
    for entry in REFERENCE_MODELS:
        rel, _, optlist = entry.partition(":")
        d = REFERENCE_ROOT / "models" / rel
        if d.is_dir():
            tasks += [(d, p, {o: True for o in optlist.split(",") if o}) for p in precisions]

    def one(t):
        d, p, opts = t
        res = [build_from_dir(str(d), precision=p, options=opts)]
        if verbose:
            print(f"built {res[0]}")
        name = res[0].name[len("libtclb_"):-len(f"_{p}.so")]
        if name in REFERENCE_ADJOINT:
            res.append(build_adjoint_lib(name, p))
        return res

    with cf.ThreadPoolExecutor(max_workers=jobs or max(1, (os.cpu_count() or 4))) as ex:
        for res in ex.map(one, tasks):
            out += res
    return out


def build_all(models: Optional[List[str]] = None, precisions=("float", "double"), force=False, verbose=False, jobs=None) -> None:
    build_core(force=force, verbose=verbose)
    models = models or all_variants()
    tasks = [(m, p) for m in models for p in precisions]
    with cf.ThreadPoolExecutor(max_workers=jobs or max(1, (os.cpu_count() or 4))) as ex:
        futs = {ex.submit(build_model_lib, m, p, force, verbose): (m, p) for m, p in tasks}
        for f in cf.as_completed(futs):
            f.result()
    # adjoint companions (dual-number build) of the models the adjoint tests and examples use
    for m in ADJOINT_MODELS:
        if m in models:
            for p in precisions:
                if p in ("float", "double"):
                    build_adjoint_lib(m, p, force=force)


ADJOINT_MODELS = ["d2q9", "d3q19"]


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("models", nargs="*", help="model variants (default: all registered)")
    ap.add_argument("--precision", choices=["float", "double", "both"], default="both")
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--list", action="store_true")
    ap.add_argument("--core-only", action="store_true")
    ap.add_argument("--dir", action="append", default=[], help="build from an unmodified TCLB model directory (repeatable)")
    ap.add_argument("--option", action="append", default=[], help="model option to switch on for --dir builds (repeatable)")
    a = ap.parse_args(argv)
    if a.dir:
        build_core(force=a.force, verbose=True)
        for d in a.dir:
            for pr in (("float", "double") if a.precision == "both" else (a.precision,)):
                print(build_from_dir(d, precision=pr, options={o: True for o in a.option}))
        return 0
    if a.list:
        print("\n".join(all_variants()))
        return 0
    if a.core_only:
        build_core(force=a.force, verbose=True)
        return 0
    prec = ("float", "double") if a.precision == "both" else (a.precision,)
    build_all(a.models or None, prec, a.force, verbose=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
