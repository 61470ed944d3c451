"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line.

usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass -k regex:K > dump.csv
       python tools/ncu_lines.py dump.csv [top_n]
"""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
fname, hdr = None, None
agg = defaultdict(lambda: [0, 0, 0, "", defaultdict(int)])  # samples, inst, thread_inst, src, stalls
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
        stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0].strip():           # a CUDA source line row
        cur = (fname, int(r[0]))
        agg[cur][3] = r[1].strip()[:100]
        continue
    if cur is None:
        continue
    try:
        s, n, t = int(r[si] or 0), int(r[ii] or 0), int(r[ti] or 0)
    except ValueError:
        continue
    a = agg[cur]
    a[0] += s
    a[1] += n
    a[2] += t
    for i, h in stall_cols:
        try:
            a[4][h] += int(r[i] or 0)
        except ValueError:
            pass
ts = sum(a[0] for a in agg.values()) or 1
tn = sum(a[1] for a in agg.values()) or 1
print(f"total samples {ts}  warp-instructions {tn}")
print(f"{'samp%':>6} {'inst%':>6} {'thr/inst':>8}  file:line  source   [top stalls]")
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    st = sorted(a[4].items(), key=lambda kv: -kv[1])[:3]
    sts = " ".join(f"{k[6:]}={100 * v // max(a[0], 1)}%" for k, v in st if v)
    print(f"{100 * a[0] / ts:6.2f} {100 * a[1] / tn:6.2f} {a[2] / max(a[1], 1):8.1f}  {key[0]}:{key[1]}  {a[3]}   [{sts}]")
