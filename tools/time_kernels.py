"""CUDA-event timings of the hot kernels on the bench workload (quick feedback loop)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import head, synth  # noqa: E402
from yolo_v1_b200.loss import loss_fwd_bwd  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--images", type=int, default=65536)
ap.add_argument("--profiles", default="uniform,realistic,correlated")
ap.add_argument("--conf", type=float, default=0.3)
a = ap.parse_args()
thr = [0.5 + 0.05 * i for i in range(10)]


def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for prof in a.profiles.split(","):
    tgt = synth.make_targets(a.images, seed=1234, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(a.images, prof, seed=1234, targets=tgt).cuda()
    tgt = tgt.cuda()
    gt = torch.zeros(80, dtype=torch.int32, device="cuda")
    ch = head.map_update(pred, tgt, a.conf, 0.4, thr, gt)
    k, f, off = head.compact_records(ch)
    m = int(off[-1].item())
    out = {
        "loss": t(lambda: loss_fwd_bwd(pred, tgt, 7, 2, 20, 5.0, 0.5, True)),
        "map_update": t(lambda: head.map_update(pred, tgt, a.conf, 0.4, thr, gt)),
        "map_update_nolists": t(lambda: head.map_update(pred, tgt, a.conf, 0.4, thr, gt, want_lists=False)),
        "decode_metrics": t(lambda: head.decode_nms(pred, a.conf, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, want_boxes=False)),
        "decode_infer": t(lambda: head.decode_nms(pred, a.conf, 0.4, head.IOU_INFER, head.ORDER_GLOBAL_CONF, want_boxes=False)),
        "compact": t(lambda: head.compact_records(ch)),
    }
    tc = t(lambda: (k[:m].clone(), f[:m].clone()))
    out["ap_reduce"] = t(lambda: head.ap_reduce(k[:m].clone(), f[:m].clone(), m, gt, 20, 10)) - tc
    print(prof, f"kept/img {m / a.images:.1f}", " ".join(f"{n}={v:.3f}" for n, v in out.items()), flush=True)
