"""Stall-reason totals of an `ncu --page source --csv --print-source cuda,sass` dump, for all lines or a line range of a file.
usage: ncu_stalls.py dump.csv [file.cu lo hi]"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
sel = (sys.argv[2], int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else None
fname = hdr = cur = None
tot = defaultdict(int)
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        cols = [(i, n) for i, n in enumerate(hdr) if n.startswith("stall_") and "Not Issued" not in n]
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0].strip():
        cur = (fname, int(r[0]))
        continue
    if cur is None or (sel and not (cur[0] == sel[0] and sel[1] <= cur[1] <= sel[2])):
        continue
    for i, n in cols:
        try:
            tot[n] += int(r[i] or 0)
        except ValueError:
            pass
s = sum(tot.values()) or 1
print("samples", s, "range", sel)
for n, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v:
        print(f"  {n:26s} {v:8d} {100 * v / s:5.1f}%")
