#!/usr/bin/env python
"""Executed warp-instructions per SOURCE LINE of a profiled kernel.

ncu's CSV export of the source page carries metrics only in the SASS view, so the SASS rows are joined (by
instruction order) with `nvdisasm -g` line markers of the same cubin.

  python tools/ncu_lines.py <prof.ncu-rep> <lib.so> <mangled-substring-of-kernel> [top]
"""
import collections
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    return [(r[isrc].strip(), int(r[iex]), int(r[ismp] or 0)) for r in rows[2:] if len(r) > iex and r[iex].isdigit()]


def line_map(lib, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
    cubin = glob.glob(os.path.join(tmp, "*.cubin"))[0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
    lines, cur, inside = [], None, False
    for l in dis:
        if l.startswith(".text."):
            inside = kernel in l
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
        if m:
            # keep the OUTERMOST user line for inlined code ("inlined at" chains list the caller afterwards)
            cur = (m.group(1), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
            lines.append(cur)
    return lines


def main():
    rep, lib, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = sass_rows(rep)
    lm = line_map(lib, kernel)
    if len(lm) != len(rows):
        print(f"# warning: {len(rows)} profiled SASS rows vs {len(lm)} disassembled instructions", file=sys.stderr)
    warps = rows[0][1]
    per = collections.Counter()
    smp = collections.Counter()
    for (src, ex, k), ln in zip(rows, lm):
        per[ln] += ex
        smp[ln] += k
    srcs = {}
    tot = sum(per.values())
    print(f"# {rep}: {tot / warps:.1f} executed instructions per warp")
    for ln, ex in per.most_common(top):
        text = ""
        if ln:
            if ln[0] not in srcs:
                try:
                    srcs[ln[0]] = open(ln[0]).read().split("\n")
                except OSError:
                    srcs[ln[0]] = []
            s = srcs[ln[0]]
            text = s[ln[1] - 1].strip()[:110] if ln[1] - 1 < len(s) else ""
        print(f"{ex / warps:8.1f} /warp  samples {smp[ln]:6d}  line {ln[1] if ln else '?':>5}: {text}")


if __name__ == "__main__":
    main()
