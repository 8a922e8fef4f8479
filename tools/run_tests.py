#!/usr/bin/env python
"""Runner for TCLB `.test` files (the reference's tools/tests.sh): each line of a test file is a command

    need <file>...          copy inputs next to the run directory
    run <command>           `$MODELBIN/main case.xml ...` is mapped to `bin/tclb [flags] <model> case.xml ...`
    csvdiff <file> [tol]    compare against the golden copy of the test directory (tools/csvdiff rule)
    pvtidiff / vtidiff <file> [tol]   compare VTK image data field by field (bin/compare)
    exists <file>, fail <command>, sha1 <file> (checked against <file>.sha1 of the golden directory)

usage: python tools/run_tests.py <model> <test directory> [--tests a.test b.test] [--cpu|--gpu] [--precision P] [--library NAME]
exit status 0 = all tests passed.
"""
import argparse
import glob
import hashlib
import os
import shlex
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_one(test_file, model, lib, flags, verbose):
    tdir = os.path.dirname(os.path.abspath(test_file))
    wd = tempfile.mkdtemp(prefix="tclb_test_")
    ok = True
    for raw in open(test_file):
        line = raw.strip()
        if not line or line.startswith("#"):
            continue
        cmd, *args = shlex.split(line)
        if cmd == "need":
            for a in args:
                for f in glob.glob(os.path.join(tdir, a)) or [os.path.join(tdir, a)]:
                    shutil.copy(f, wd)
        elif cmd in ("run", "fail"):
            argv = []
            for a in args:
                if a.endswith("/main") or a == "$SOLVER":
                    argv += [os.path.join(ROOT, "bin", "tclb")] + flags + [lib]
                elif a.startswith("mpirun"):
                    continue
                else:
                    argv.append(a)
            # `mpirun -np N $MODELBIN/main` -> --ranks N
            if "-np" in args:
                n = args[args.index("-np") + 1]
                argv = [x for x in argv if x not in ("-np", n)]
                argv.insert(1, "--ranks"); argv.insert(2, n)
            r = subprocess.run(argv, cwd=wd, capture_output=True, text=True)
            good = (r.returncode == 0) if cmd == "run" else (r.returncode != 0)
            if not good:
                print(f"  {cmd} FAILED: {' '.join(argv)}\n{r.stdout[-1500:]}{r.stderr[-1500:]}")
                ok = False
                break
        elif cmd == "csvconcatenate":
            r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "csvconcatenate")] + args, cwd=wd, capture_output=True, text=True)
            if r.returncode != 0:
                print(f"  csvconcatenate: {r.stdout}{r.stderr}")
                ok = False
        elif cmd == "csvdiff":
            tol, gold = "1e-10", args[0]
            for extra in args[1:]:
                try:
                    float(extra); tol = extra
                except ValueError:
                    gold = extra
            r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "csvdiff"), "-a", os.path.join(wd, args[0]), "-b",
                                os.path.join(tdir, gold), "-x", tol], capture_output=True, text=True)
            if r.returncode != 0:
                print(f"  csvdiff {args[0]}: {r.stdout.strip()}")
                ok = False
        elif cmd in ("pvtidiff", "vtidiff"):
            gold = os.path.join(tdir, args[1] if len(args) > 1 and not args[1].replace(".", "").replace("e-", "").isdigit() else args[0])
            pieces = glob.glob(os.path.splitext(gold)[0] + "*.vti")
            if cmd == "pvtidiff" and not pieces:
                print(f"  {cmd} {args[0]}: golden pieces are not shipped, only checking that the file was written")
                ok = ok and os.path.exists(os.path.join(wd, args[0]))
                continue
            r = subprocess.run([os.path.join(ROOT, "bin", "compare"), os.path.join(wd, args[0]), gold], capture_output=True, text=True)
            if r.returncode != 0:
                print(f"  {cmd} {args[0]}: {r.stdout[-500:]}")
                ok = False
        elif cmd == "exists":
            if not os.path.exists(os.path.join(wd, args[0])):
                print(f"  missing {args[0]}")
                ok = False
        elif cmd == "sha1":
            want = open(os.path.join(tdir, args[0] + ".sha1")).read().split()[0]
            got = hashlib.sha1(open(os.path.join(wd, args[0]), "rb").read()).hexdigest()
            if want != got:
                print(f"  sha1 {args[0]} differs")
                ok = False
        else:
            print(f"  unknown test command '{cmd}'")
            ok = False
    if verbose or not ok:
        print(f"  work directory: {wd}")
    else:
        shutil.rmtree(wd, ignore_errors=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("model")
    ap.add_argument("directory")
    ap.add_argument("--tests", nargs="*")
    ap.add_argument("--library", default=None, help="library name if it differs from the model name (e.g. ref_d2q9)")
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--precision", default="double")
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args()
    tests = [os.path.join(a.directory, t) for t in a.tests] if a.tests else sorted(glob.glob(os.path.join(a.directory, "*.test")))
    flags = ["--gpu" if a.gpu else "--cpu", "--precision", a.precision]
    bad = 0
    for t in tests:
        ok = run_one(t, a.model, a.library or a.model, flags, a.verbose)
        print(f"{'OK  ' if ok else 'FAIL'} {os.path.basename(t)}")
        bad += 0 if ok else 1
    print(f"{len(tests) - bad} of {len(tests)} tests passed")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
