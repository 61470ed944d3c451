"""Kernel timeline of the captured head step (pipelined CUDA graphs) on every rank: which kernel each rank waits in.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/step_timeline.py --out profiles/r02_timeline_Ngpu.md

torch.profiler (CUPTI) records the kernels of 8 pipelined steps; per kernel name the mean duration over the steps is taken on
every rank and the table lists min / mean / max over the ranks next to the same kernel's duration on ONE GPU running alone
(no peers to wait for).  A kernel whose duration grows with N is one that waits: the exchange's flag waits sit inside
xi_scatter (the owners' base offsets), peer_finish (everybody's records have landed) and peer_sum_results (the AP sums)."""
import argparse
import json
import os
import re
import sys
from collections import defaultdict

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import CapturedHeadStep, YOLOLoss, mAPMetric, synth  # noqa: E402
from yolo_v1_b200.distributed import DistributedMAP  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--images", type=int, default=65536)
ap.add_argument("--steps", type=int, default=8)
ap.add_argument("--out", default="")
args = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
tgt = synth.make_targets(args.images, seed=1234 + rank, p_obj=0.055, size_mix=True)
pred = synth.make_preds(args.images, "uniform", seed=1234 + rank, targets=tgt).to(dev)
tgt = tgt.to(dev)
m = mAPMetric(num_classes=20, conf_threshold=0.3, nms_threshold=0.4)
metric = DistributedMAP(m, mode="peer") if world > 1 else m
step = CapturedHeadStep(YOLOLoss(), metric, pred, tgt, pipelined=True)


def pipelined(k):
    step.submit()
    for i in range(k):
        if i + 1 < k:
            step.submit()
        step.collect()


pipelined(10)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
pipelined(50)
e1.record()
torch.cuda.synchronize()
ms_step = e0.elapsed_time(e1) / 50
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    pipelined(args.steps)
    torch.cuda.synchronize()
path = f"/tmp/trace_{rank}.json"
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
os.remove(path)
ev.sort(key=lambda e: e["ts"])
dur = defaultdict(list)
for e in ev:
    name = re.sub(r"\(.*", "", e["name"]).replace("void ", "")
    name = re.sub(r"^at::native::.*", "torch elementwise / cat", name)
    dur[name].append(e["dur"])
mine = {k: (sum(v) / len(v), len(v) / args.steps) for k, v in dur.items()}
alls = [None] * world
if world > 1:
    dist.all_gather_object(alls, (mine, ms_step))
else:
    alls = [(mine, ms_step)]
if rank == 0:
    lines = [f"# Kernel timeline of the pipelined head step, {world} GPU(s), {args.images} uniform images per GPU",
             "", f"ms per step (CUDA events, 50 steps): " + ", ".join(f"rank {r}: {a[1]:.3f}" for r, a in enumerate(alls)), "",
             "Mean kernel duration in us over the profiled steps (torch.profiler / CUPTI), launches per step in brackets; min / mean / max over the ranks.",
             "", "| kernel | per step | min | mean | max | busiest rank |", "|---|---|---|---|---|---|"]
    names = sorted({k for a in alls for k in a[0]}, key=lambda k: -max(a[0].get(k, (0, 0))[0] * a[0].get(k, (0, 0))[1] for a in alls))
    tot = [0.0] * world
    for k in names:
        vals = [a[0].get(k, (0.0, 0.0)) for a in alls]
        per = [v[0] * v[1] for v in vals]           # us per step
        for r in range(world):
            tot[r] += per[r]
        if max(per) < 1.0:
            continue
        lines.append(f"| `{k[:60]}` | x{max(v[1] for v in vals):.1f} | {min(per):.1f} | {sum(per) / world:.1f} | {max(per):.1f} | {per.index(max(per))} |")
    lines += ["", "Sum of kernel time per step (us; kernels of the two streams overlap, so this is more than the step): " +
              ", ".join(f"rank {r}: {t:.0f}" for r, t in enumerate(tot)), ""]
    # one step of rank 0 in order
    lines += ["## Rank 0, one step in launch order (start us relative to the update kernel, duration us, stream)", "", "```"]
    starts = [i for i, e in enumerate(ev) if "map_update_kernel" in e["name"]]
    if len(starts) >= 4:
        lo, hi = starts[2], starts[3]
        t0 = ev[lo]["ts"]
        for e in ev[lo:hi]:
            lines.append(f"{e['ts'] - t0:9.1f} {e['dur']:8.1f} s{e['args'].get('stream')} {re.sub(r'^void ', '', e['name'])[:80]}")
    lines.append("```")
    text = "\n".join(lines) + "\n"
    if args.out:
        open(args.out, "w").write(text)
    print(text)
if world > 1:
    metric.close()
    dist.destroy_process_group()
