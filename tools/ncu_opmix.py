#!/usr/bin/env python
"""Executed-instruction mix of the profiled kernel, from `ncu --page source --csv` (SASS view).

  python tools/ncu_opmix.py gpurun_out/prof.ncu-rep [top]
"""
import collections
import csv
import io
import subprocess
import sys


def load(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    isrc, iex = hdr.index("Source"), hdr.index("Instructions Executed")
    ismp = hdr.index("# Samples")
    res = []
    for r in rows[2:]:
        if len(r) <= iex:
            continue
        try:
            res.append((r[isrc].strip(), int(r[iex]), int(r[ismp] or 0)))
        except ValueError:
            pass
    return res


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    ins = load(rep)
    warps = ins[0][1]            # the entry instruction runs once per warp
    mix = collections.Counter()
    smp = collections.Counter()
    for s, e, k in ins:
        t = s.split()
        op = t[1] if t and t[0].startswith("@") else (t[0] if t else "?")
        op = op.split(".")[0]
        mix[op] += e
        smp[op] += k
    tot = sum(mix.values())
    print(f"# {rep}: {tot} warp instructions, {warps} warps -> {tot / warps:.1f} executed instructions per warp")
    for op, e in mix.most_common(top):
        print(f"{op:12s} {e / warps:8.1f} /warp  {100.0 * e / tot:5.1f}%   samples {smp[op]}")


if __name__ == "__main__":
    main()
