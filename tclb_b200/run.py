"""``python -m tclb_b200.run [--precision P] [--cpu|--gpu] [--ranks N] <model> <case.xml> [XPATH ...]``

Same as ``bin/tclb`` plus the in-process Python coupling: ``<RunPython>`` elements of the case run in this interpreter
(see :mod:`tclb_b200.solver`)."""
import argparse
import sys

from .solver import Solver


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m tclb_b200.run")
    ap.add_argument("--precision", default="double", choices=["float", "double"])
    g = ap.add_mutually_exclusive_group()
    g.add_argument("--cpu", action="store_true")
    g.add_argument("--gpu", action="store_true")
    ap.add_argument("--ranks", type=int, default=1)
    ap.add_argument("model")
    ap.add_argument("case")
    ap.add_argument("xpath", nargs="*")
    a = ap.parse_args(argv)
    s = Solver(a.model, a.precision, "cuda" if a.gpu else "cpu", nranks=a.ranks)
    s.load(a.case, a.xpath).run()
    s.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
