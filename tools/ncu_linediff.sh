#!/bin/bash
# usage: ncu_linediff.sh A.ncu-rep kernelA B.ncu-rep kernelB lib.so [top]   -- per-source-line executed-instruction difference A-B
python tools/ncu_lines.py $1 $5 $2 4000 > /tmp/_a.txt; python tools/ncu_lines.py $3 $5 $4 4000 > /tmp/_b.txt
head -1 /tmp/_a.txt /tmp/_b.txt
python - ${6:-25} <<'PY'
import re,sys
def rd(p):
    d={}; t={}
    for l in open(p):
        m=re.match(r"\s*([\d.]+) /warp\s+samples\s+(\d+)\s+line\s+(\S+): (.*)",l)
        if m: d[m.group(3)]=float(m.group(1)); t[m.group(3)]=m.group(4)
    return d,t
g,tg=rd('/tmp/_a.txt'); n,tn=rd('/tmp/_b.txt')
tg.update(tn)
ks=sorted(set(g)|set(n), key=lambda k:-(g.get(k,0)-n.get(k,0)))
for k in ks[:int(sys.argv[1])]: print(f"{g.get(k,0):7.1f} {n.get(k,0):7.1f} {g.get(k,0)-n.get(k,0):7.1f}  {k}: {tg[k][:100]}")
PY
