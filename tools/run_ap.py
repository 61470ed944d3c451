"""compact + ap_reduce a few times on the bench workload (for ncu captures of the reduce-side kernels)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import mAPMetric, synth
n = 65536
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, os.environ.get("PROFILE", "uniform"), seed=1234, targets=tgt).cuda(); tgt = tgt.cuda()
m = mAPMetric(num_classes=20, conf_threshold=0.3)
for _ in range(3):
    m.reset(); m.update(pred, tgt); m.compute()
torch.cuda.synchronize()
