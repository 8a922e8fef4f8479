#!/usr/bin/env python
"""Executed-instruction mix per kernel of an .ncu-rep (from the SASS source page):  python tools/ncu_opmix.py rep [kernel-name substring]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]
raw = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)))
h = raw[0]
names = [r[h.index("Kernel Name")] for r in raw[2:]]
grid = [int(r[h.index("launch__grid_size")]) for r in raw[2:]]
block = [int(r[h.index("launch__block_size")]) for r in raw[2:]]
dur = [r[h.index("gpu__time_duration.sum")] for r in raw[2:]]
for k, name in enumerate(names):
    if len(sys.argv) > 2 and sys.argv[2] not in name:
        continue
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-skip", str(k), "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hh = rows[1]
    iS, iE = hh.index("Source"), hh.index("Instructions Executed")
    hist, tot = collections.Counter(), 0
    for r in rows[2:]:
        try:
            n = int(r[iE])
        except (ValueError, IndexError):
            continue
        s = r[iS].split()
        if not s:
            continue
        op = (s[1] if s[0].startswith("@") else s[0]).split(".")[0]
        hist[op] += n
        tot += n
    print(f"[{k}] {name[:70]}  {dur[k]} ms  total {tot} warp-instr")
    print("    " + " ".join(f"{o}:{v * 100.0 / max(tot, 1):.1f}%" for o, v in hist.most_common(22)))
