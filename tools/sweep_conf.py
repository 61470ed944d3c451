"""decode_nms / map_update time against the amount of per-image work (confidence threshold sweep)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import head, synth  # noqa: E402

n = 65536
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, "uniform", seed=1234, targets=tgt).cuda()
tgt = tgt.cuda()
zero_t = torch.zeros_like(tgt)
thr = [0.5 + 0.05 * i for i in range(10)]
gt = torch.zeros(80, dtype=torch.int32, device="cuda")


def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for conf in (0.3, 0.5, 0.7, 0.9, 0.99, 2.0):
    r = head.decode_nms(pred, conf, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, want_boxes=False)
    print(f"conf {conf}: cand/img {r.n_cand.float().mean().item():5.1f} kept/img {r.n_keep.float().mean().item():5.1f}  "
          f"decode_metrics {t(lambda: head.decode_nms(pred, conf, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, want_boxes=False)):.3f}  "
          f"decode_infer {t(lambda: head.decode_nms(pred, conf, 0.4, head.IOU_INFER, head.ORDER_GLOBAL_CONF, want_boxes=False)):.3f}  "
          f"map_update {t(lambda: head.map_update(pred, tgt, conf, 0.4, thr, gt)):.3f}  "
          f"map_update(no gt) {t(lambda: head.map_update(pred, zero_t, conf, 0.4, thr, gt)):.3f}", flush=True)
