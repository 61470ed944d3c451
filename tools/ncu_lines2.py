"""Per source line: warp instructions and stall samples from an `ncu --page source --csv --print-source cuda,sass`
dump.  usage: ncu_lines2.py dump.csv [top_n]   (lines sorted by file/line; share of the kernel's totals)"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 0
fname = hdr = None
agg = defaultdict(lambda: [0, 0, 0, ""])
cur = None
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
        agg[cur][3] = r[1].strip()[:110]
        continue
    if cur is None:
        continue
    try:
        agg[cur][0] += int(r[si] or 0); agg[cur][1] += int(r[ii] or 0); agg[cur][2] += int(r[ti] or 0)
    except ValueError:
        pass
ts = sum(v[0] for v in agg.values()) or 1
tn = sum(v[1] for v in agg.values()) or 1
print(f"total warp instructions {tn}, samples {ts}")
items = sorted(agg.items(), key=(lambda kv: -kv[1][1]) if top else (lambda kv: kv[0]))
for (f, ln), (s, i, t, src) in (items[:top] if top else items):
    if i * 1000 < tn and s * 1000 < ts:
        continue
    print(f"{f:16s}{ln:5d} inst {100*i/tn:5.2f}% samp {100*s/ts:5.2f}% thr {t/max(i,1):4.1f} | {src}")
