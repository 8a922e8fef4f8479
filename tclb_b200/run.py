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
    ap.add_argument("--halo", default="direct", choices=["direct", "nccl"], help="multi-process runs: fused peer stores or the NCCL/gloo baseline")
    a = ap.parse_args(argv)
    import os
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        return run_distributed(a, int(os.environ.get("RANK", "0")), world, int(os.environ.get("LOCAL_RANK", "0")))
    s = Solver(a.model, a.precision, "cuda" if a.gpu else "cpu", nranks=a.ranks)
    s.load(a.case, a.xpath).run()
    s.close()
    return 0


def run_distributed(a, rank: int, world: int, local_rank: int) -> int:
    """One process per GPU (``python -m torch.distributed.run --nproc-per-node N -m tclb_b200.run --gpu model case.xml``):
    every process owns ONE subdomain of the reference's y/z decomposition and runs the same handler tree.  What the
    reference does with MPI calls inside its handlers happens here as follows: the halo and the Globals travel inside the
    stage kernels (peer memory), so every rank holds the same Globals and takes the same Stop/PID/Keep decisions; rank 0
    writes the Log, the `.pvti` index and the restart XML, every rank its own `.vti` piece and `.pri` file
    (src/vtkOutput.cpp:134-148, src/Handlers/cbSaveCheckpoint.cpp); Failcheck reduces its flag through the process group
    (src/Handlers/cbFailcheck.cpp:74)."""
    from .parallel.dist import DistDomain
    from .runtime import core
    backend = "cuda" if a.gpu else "cpu"
    core().tclb_log_init(rank, 3)          # L_OUTPUT; messages below level 6 print on rank 0 only
    s = Solver(a.model, a.precision, backend, nranks=world, local_ranks=[rank], devices=[local_rank])
    s.load(a.case, a.xpath)
    dd = DistDomain(s.model, None, rank=rank, world=world, device=local_rank, backend=backend, direct=(a.halo == "direct"), domain=s.domain)
    try:
        s.run()
    finally:
        s.domain.sync()
        dd.barrier()
    s.close()
    if dd.dist.is_initialized():
        dd.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
