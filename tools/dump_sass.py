#!/usr/bin/env python
"""Write one SASS listing per kernel under profiles/sass/<library>/ plus an index with the mnemonic counts that show
which data-movement instructions a kernel uses (UBLKCP = cp.async.bulk, SYNCS = mbarrier, LDGSTS = cp.async, ...).

    python tools/dump_sass.py                 # the benchmarked libraries + the core
    python tools/dump_sass.py libtclb_d2q9_float.so --all-modes

Encodings are dropped from the listing (the instruction text is what is read); only the Globals modes the benchmarks
run (0 = none, 1 = integrate) are kept unless --all-modes is given.
"""
import argparse
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "tclb_b200", "_lib")
DEFAULT = ["libtclb_core.so", "libtclb_d3q27_cumulant_float.so", "libtclb_d3q19_float.so", "libtclb_ref_d3q27_pf_velocity_float.so",
           "libtclb_ref_d3q27_cumulant_float.so", "libtclb_ref_d3q19_float.so", "libtclb_d2q9_float_ad.so"]
WATCH = ["UBLKCP", "SYNCS", "LDGSTS", "UTMALDG", "LDG", "STG", "LDS", "STS", "LDL", "STL", "ATOM", "ATOMG", "RED", "FFMA", "FADD", "FMUL", "DFMA",
         "MUFU", "BAR", "BRA", "SHFL", "NANOSLEEP"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def short(d):
    d = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", d)
    d = re.sub(r"\(.*$", "", d)
    d = re.sub(r"^void ", "", d)
    d = re.sub(r"\(int\)|\(bool\)", "", d)
    return re.sub(r"[^A-Za-z0-9_<>,-]", "", d).replace("<", "_").replace(">", "").replace(",", "_")


def keep(name, all_modes):
    m = re.match(r"k_stage(_persist)?_(\d+)_(\d+)", name)
    if m and not all_modes and int(m.group(3)) not in (0, 1):
        return False
    m = re.match(r"k_fold_globals_(\d+)", name)
    if m and not all_modes and int(m.group(1)) not in (1,):
        return False
    return True


def dump(lib, outroot, all_modes):
    path = lib if os.path.exists(lib) else os.path.join(LIBDIR, lib)
    if not os.path.exists(path):
        print("skip (not built):", lib)
        return []
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    names = [p.split("\n", 1)[0].strip() for p in parts[1:]]
    dm = demangle(names)
    label = os.path.basename(path).replace("libtclb_", "").replace(".so", "")
    odir = os.path.join(outroot, label)
    os.makedirs(odir, exist_ok=True)
    rows = []
    for p, mangled in zip(parts[1:], names):
        name = short(dm[mangled])
        if not keep(name, all_modes):
            continue
        lines, hist = [], collections.Counter()
        for ln in p.split("\n")[1:]:
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?)\s*/\* 0x[0-9a-f]+ \*/", ln)
            if m:
                ins = m.group(2).rstrip(" ;")
                lines.append(f"/*{m.group(1)}*/ {ins}")
                op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
                hist[op.split(".")[0]] += 1
            elif ".headerflags" in ln or "Function" in ln:
                continue
        with open(os.path.join(odir, name + ".sass"), "w") as f:
            f.write(f"// {dm[mangled]}\n// {os.path.basename(path)}, cuobjdump -sass, encodings dropped\n" + "\n".join(lines) + "\n")
        rows.append((label, name, len(lines), hist))
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("libs", nargs="*")
    ap.add_argument("--all-modes", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "sass"))
    a = ap.parse_args()
    rows = []
    for lib in a.libs or DEFAULT:
        rows += dump(lib, a.out, a.all_modes)
    with open(os.path.join(a.out, "INDEX.md"), "w") as f:
        f.write("# SASS listings (sm_100a), one file per kernel\n\nStatic instruction counts per mnemonic; `UBLKCP` = `cp.async.bulk` (the TMA unit's bulk copy), "
                "`SYNCS` = mbarrier operations, `LDGSTS` = `cp.async`.\nRegenerate with `python tools/dump_sass.py`.\n\n")
        f.write("| library | kernel | instructions | " + " | ".join(WATCH) + " |\n|---|---|---|" + "---|" * len(WATCH) + "\n")
        for label, name, n, hist in rows:
            f.write(f"| {label} | [{name}]({label}/{name}.sass) | {n} | " + " | ".join(str(hist.get(w, 0)) for w in WATCH) + " |\n")
    print(f"{len(rows)} kernels -> {a.out}")


if __name__ == "__main__":
    sys.exit(main())
