"""Per-kernel GPU times of one bench step on one GPU (torch profiler / CUPTI), averaged over a few steps."""
import os, sys
import torch
from torch.profiler import ProfilerActivity, profile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import YOLOLoss, mAPMetric, synth

n = int(os.environ.get("N", 65536))
prof_name = os.environ.get("PROFILE", "uniform")
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, prof_name, seed=1234, targets=tgt).cuda()
tgt = tgt.cuda()
crit = YOLOLoss(defer_loss_dict=True)
metric = mAPMetric(num_classes=20, conf_threshold=float(os.environ.get("CONF", 0.3)))


def step():
    p = pred.detach().requires_grad_(True)
    total, parts = crit(p, tgt)
    total.backward()
    metric.reset()
    metric.update(p.detach(), tgt)
    return float(parts["total"]), metric.compute()["mAP50:95"]


for _ in range(5):
    step()
torch.cuda.synchronize()
reps = 5
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(reps):
        step()
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / reps, e.count / reps) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows)
print(f"{'kernel':70s} {'us/step':>9s} {'calls':>6s}")
for k, us, c in rows:
    print(f"{k[:70]:70s} {us:9.1f} {c:6.1f}")
print(f"{'sum':70s} {tot:9.1f}")
