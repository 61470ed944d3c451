"""Wall-clock split of one bench step (device-resident inputs): where the non-kernel time goes."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import YOLOLoss, mAPMetric, synth

n = 65536
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, "uniform", seed=1234, targets=tgt).cuda()
tgt = tgt.cuda()
crit = YOLOLoss()
metric = mAPMetric(num_classes=20, conf_threshold=0.3)


def step(sync):
    t = [time.perf_counter()]
    def mark():
        if sync:
            torch.cuda.synchronize()
        t.append(time.perf_counter())
    p = pred.detach().requires_grad_(True)
    total, parts = crit(p, tgt); mark()
    total.backward(); mark()
    metric.reset(); metric.update(p.detach(), tgt); mark()
    res = metric.compute(); mark()
    return [1e3 * (b - a) for a, b in zip(t, t[1:])]


for _ in range(5):
    step(False)
for sync in (True, False):
    acc = None
    for _ in range(20):
        torch.cuda.synchronize()
        d = step(sync)
        acc = d if acc is None else [a + b for a, b in zip(acc, d)]
    print("sync" if sync else "async", "loss_fwd %.3f backward %.3f update %.3f compute %.3f  total %.3f ms" % (*[a / 20 for a in acc], sum(acc) / 20))

# finer: host time spent after the packed result is back (dict assembly), and enqueue latency of the first kernel
import yolo_v1_b200.metrics as M
orig = metric._results
spent = []
def timed(*a, **k):
    t0 = time.perf_counter(); r = orig(*a, **k); spent.append(time.perf_counter() - t0); return r
metric._results = timed
crit2 = YOLOLoss(defer_loss_dict=True)
first = []
for _ in range(20):
    torch.cuda.synchronize()
    p = pred.detach().requires_grad_(True)
    t0 = time.perf_counter(); total, parts = crit2(p, tgt); first.append(time.perf_counter() - t0)
    total.backward(); metric.reset(); metric.update(p.detach(), tgt); metric.compute(); float(parts["total"])
print("results assembly %.3f ms   loss forward enqueue %.3f ms" % (1e3 * sum(spent) / len(spent), 1e3 * sum(first) / len(first)))
