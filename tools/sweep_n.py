"""Per-image time of the per-image kernels against the batch size: small batches stay in the 126 MB L2 across
repeats (no DRAM traffic), large ones stream from HBM.  Separates the compute floor from the memory floor."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import head, synth  # noqa: E402

thr = [0.5 + 0.05 * i for i in range(10)]
gt = torch.zeros(80, dtype=torch.int32, device="cuda")


def t(fn, reps=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


tgt_all = synth.make_targets(65536, seed=1234, p_obj=0.055, size_mix=True)
pred_all = synth.make_preds(65536, "uniform", seed=1234, targets=tgt_all)
for n in (4736, 9472, 18944, 37888, 65536):
    pred, tgt = pred_all[:n].cuda(), tgt_all[:n].cuda()
    for conf in (0.3, 2.0):
        a = t(lambda: head.decode_nms(pred, conf, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, want_boxes=False))
        b = t(lambda: head.map_update(pred, tgt, conf, 0.4, thr, gt))
        print(f"n {n:6d} ({n * 11760 / 1e6:5.0f} MB) conf {conf}: decode_metrics {a * 1e6 / n:6.2f} ns/img ({a:.3f} ms)   "
              f"map_update {b * 1e6 / n:6.2f} ns/img ({b:.3f} ms)", flush=True)
