"""Every kernel of the library once at small sizes, for compute-sanitizer (memcheck / racecheck / initcheck).

    compute-sanitizer --tool memcheck python tools/sanitize.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import (CapturedHeadStep, YOLOInference, YOLOLoss, encode_targets, head, mAPMetric,  # noqa: E402
                          synth)
from yolo_v1_b200 import distributed as D  # noqa: E402

dev = torch.device("cuda")
n = int(os.environ.get("N", 333))
tgt = synth.make_targets(n, seed=3, p_obj=0.08, size_mix=True, both_slots=0.1)
for profile, conf in (("correlated", 0.01), ("uniform", 0.3), ("ties", 0.01)):
    pred = synth.make_preds(n, profile, seed=4, targets=tgt)
    p = pred.to(dev).requires_grad_(True)
    total, parts = YOLOLoss()(p, tgt.to(dev))
    (total * 3.0).backward()
    m = mAPMetric(num_classes=20, conf_threshold=conf)
    m.update(pred[:100].to(dev), tgt[:100].to(dev))
    m.update(pred[100:].to(dev), tgt[100:].to(dev))
    res = m.compute()
    _ = m.all_predictions, m.all_ground_truths            # list forms
    m2 = mAPMetric(num_classes=20, conf_threshold=conf)
    m2.all_predictions = m.all_predictions[:50]
    m2.all_ground_truths = m.all_ground_truths[:50]
    m2.compute()                                          # match_lists path
    head.decode_nms(pred.to(dev), conf, 0.4, head.IOU_INFER, head.ORDER_GLOBAL_CONF)
    head.decode_nms(pred.to(dev), conf, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED)
    print(profile, float(total), res["mAP50:95"])
# half precision loss
for dt in (torch.float16, torch.bfloat16):
    p = synth.make_preds(n, "uniform", seed=5).to(dev).to(dt).requires_grad_(True)
    YOLOLoss()(p, tgt.to(dev))[0].backward()
# encoder
encode_targets([[[0.5, 0.5, 0.2, 0.2], [0.1, 0.9, 0.3, 0.1]], [], [[0.99, 0.01, 0.5, 0.5]]], [[1, 2], [], [19]], device=dev)
# partition + the peer exchange with one rank (self-mapped block), graph capture
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29533")
import torch.distributed as dist  # noqa: E402

dist.init_process_group("gloo", rank=0, world_size=1)
pred = synth.make_preds(n, "correlated", seed=4, targets=tgt).to(dev)
for mode in ("class_shard", "peer"):
    dm = D.DistributedMAP(mAPMetric(num_classes=20), mode=mode)
    dm.update(pred, tgt.to(dev))
    print(mode, dm.compute()["mAP50:95"])
    dm.close()
k = torch.randint(0, 2**62, (5000,), device=dev)
f = (torch.randint(0, 20, (5000,), device=dev) << 56)
D.partition_records(k, f, 5000, D.class_owners(20, 4), 4)
step = CapturedHeadStep(YOLOLoss(), mAPMetric(num_classes=20), pred.clone(), tgt.to(dev).clone())
print("graph", step.replay()[1]["mAP50:95"])
dist.destroy_process_group()
torch.cuda.synchronize()
print("done")
