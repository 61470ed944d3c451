"""cProfile of the host side of one bench step (device-resident inputs): which Python calls cost what."""
import cProfile, os, pstats, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import YOLOLoss, mAPMetric, synth

n = int(os.environ.get("N", 65536))
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, "uniform", seed=1234, targets=tgt).cuda()
tgt = tgt.cuda()
crit = YOLOLoss(defer_loss_dict=True)
metric = mAPMetric(num_classes=20, conf_threshold=0.3)


def step():
    p = pred.detach().requires_grad_(True)
    total, parts = crit(p, tgt)
    total.backward()
    metric.reset()
    metric.update(p.detach(), tgt)
    res = metric.compute()
    return float(parts["total"]), res["mAP50:95"]


for _ in range(10):
    step()
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
