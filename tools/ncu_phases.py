"""Bucket an ncu source-page dump by (file, line range).  usage: ncu_phases.py dump.csv 'name:file:lo:hi' ..."""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
fname = hdr = cur = None
agg = defaultdict(lambda: [0, 0, 0])
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        si, ii, ti = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0].strip():
        cur = (fname, int(r[0]))
        continue
    if cur is None:
        continue
    try:
        agg[cur][0] += int(r[si] or 0); agg[cur][1] += int(r[ii] or 0); agg[cur][2] += int(r[ti] or 0)
    except ValueError:
        pass
ts = sum(v[0] for v in agg.values()) or 1
tn = sum(v[1] for v in agg.values()) or 1
used = set()
for spec in sys.argv[2:]:
    name, f, a, b = spec.split(":")
    ks = [k for k in agg if k[0] == f and int(a) <= k[1] < int(b)]
    used |= set(ks)
    s = sum(agg[k][0] for k in ks); i = sum(agg[k][1] for k in ks); t = sum(agg[k][2] for k in ks)
    print(f"{name:16s} samp {100*s/ts:5.1f}%  inst {100*i/tn:5.1f}%  thr/inst {t/max(i,1):5.1f}")
rest = sorted(((agg[k][1], k) for k in agg if k not in used), reverse=True)
print(f"rest inst {100*sum(v for v,_ in rest)/tn:.1f}%", [(k, round(100*v/tn, 2)) for v, k in rest[:10]])
