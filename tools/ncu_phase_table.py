"""Phase table of a per-image kernel from an `ncu --page source --csv --print-source cuda,sass` dump:
share of stall samples, warp instructions, shared-memory wavefronts and active threads per phase.
usage: ncu_phase_table.py dump.csv   (phases are source-line ranges of the current csrc/ files)"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
fname = hdr = cur = None
agg = defaultdict(lambda: [0, 0, 0, 0])
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        si, ii, ti, wi = (hdr.index(x) for x in ("# Samples", "Instructions Executed", "Thread Instructions Executed",
                                                  "L1 Wavefronts Shared"))
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0].strip():
        cur = (fname, int(r[0]))
        continue
    try:
        for j, col in enumerate((si, ii, ti, wi)):
            agg[cur][j] += int(r[col] or 0)
    except ValueError:
        pass


def grep_line(path, needle, after=0):
    for i, l in enumerate(open(path), 1):
        if needle in l:
            return i + after
    raise SystemExit(f"{needle!r} not in {path}")


C = "yolo_v1_b200/csrc/"
hi, hn, mu, co, fi = C + "head_image.cuh", C + "head_nms.cuh", C + "map_update.cu", C + "common.cuh", C + "iou_filter.cuh"
L = grep_line
phases = [
    ("staging", "head_image.cuh", L(hi, "void stage_half("), L(hi, "// Pull one image")),
    ("decode", "head_image.cuh", L(hi, "Mask128 decode_image("), 10 ** 6),
    ("decode: class max", "common.cuh", L(co, "int cell_class_shared("), L(co, "int cell_class_global(")),
    ("decode: fp32 box", "iou_filter.cuh", L(fi, "abs_coord_f("), L(fi, "int iou_below_fast(")),
    ("class members", "head_nms.cuh", L(hn, "int build_members("), L(hn, "void rank_members(")),
    ("pair loop", "head_nms.cuh", L(hn, "Mask128 order_and_nms("), L(hn, "for (int p = lane; p < n; p += 32) {\n") if False else L(hn, "const int r = lo + no.beaten[p];") - 3),
    ("pair: box loads + verdict", "head_nms.cuh", L(hn, "BoxF cand_box("), L(hn, "int build_members(")),
    ("pair: fp32 interval test", "iou_filter.cuh", L(fi, "int iou_below_fast("), L(fi, "YH_HD int iou_interval(")),
    ("rank scatter + greedy scan", "head_nms.cuh", L(hn, "const int r = lo + no.beaten[p];") - 3, L(hn, "void publish_keep(")),
    ("keep words", "head_nms.cuh", L(hn, "void publish_keep("), 10 ** 6),
]
if "map_update" in open(sys.argv[1]).read(4000):
    phases += [
        ("target parse", "map_update.cu", L(mu, "int parse_targets("), L(mu, "int match_exact(")),
        ("image loop head", "map_update.cu", L(mu, "for (int img = blockIdx.x"), L(mu, "// ---- targets: parse into the overlay")),
        ("records + items", "map_update.cu", L(mu, "// ---- records, in sorted order among the kept"), L(mu, "// ---- IoU intervals of all")),
        ("pair intervals", "map_update.cu", L(mu, "// ---- IoU intervals of all"), L(mu, "// ---- best ground truth and threshold mask")),
        ("best GT + masks", "map_update.cu", L(mu, "// ---- best ground truth and threshold mask"), L(mu, "// ---- greedy TP/FP")),
        ("interval arithmetic (pairs)", "iou_filter.cuh", L(fi, "YH_HD int iou_interval("), 10 ** 6),
        ("TP/FP + patch records", "map_update.cu", L(mu, "// ---- greedy TP/FP"), L(mu, "// keep the (slow) ticket read down here")),
        ("next ticket + tail", "map_update.cu", L(mu, "// keep the (slow) ticket read down here"), 10 ** 6),
    ]
ts, tn, tw = (sum(v[j] for v in agg.values()) or 1 for j in (0, 1, 3))
used = set()
print(f"| phase | samples % | warp instr % | instr / image | smem wavefronts % | threads / instr |\n|---|---|---|---|---|---|")
for name, f, a, b in phases:
    ks = [k for k in agg if k[0] == f and a <= k[1] < b and k not in used]
    used |= set(ks)
    s, i, t, w = (sum(agg[k][j] for k in ks) for j in range(4))
    print(f"| {name} | {100*s/ts:.1f} | {100*i/tn:.1f} | {i/65536:.0f} | {100*w/tw:.1f} | {t/max(i,1):.1f} |")
rest = [k for k in agg if k not in used]
s, i, t, w = (sum(agg[k][j] for k in rest) for j in range(4))
print(f"| everything else (intrinsics headers, atomics, syncs) | {100*s/ts:.1f} | {100*i/tn:.1f} | {i/65536:.0f} | {100*w/tw:.1f} | {t/max(i,1):.1f} |")
print(f"\ntotal: {tn/65536:.0f} warp instructions and {tw/65536:.0f} shared-memory wavefronts per image")
