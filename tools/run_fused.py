"""Launch the fused loss + update kernel a few times (for ncu) and time it against empty targets."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import head, synth  # noqa: E402

thr = [0.5 + 0.05 * i for i in range(10)]
tgt = synth.make_targets(65536, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(65536, "uniform", seed=1234, targets=tgt).cuda()
tgt = tgt.cuda()
gt = torch.zeros(80, dtype=torch.int32, device="cuda")


def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for _ in range(3):
    head.map_update_loss(pred, tgt, 0.3, 0.4, thr, gt, 5.0, 0.5, True)
torch.cuda.synchronize()
if "--time" in sys.argv:
    zero = torch.zeros_like(tgt)
    for name, tg in (("targets", tgt), ("empty targets", zero)):
        print(f"{name}: map_update {t(lambda: head.map_update(pred, tg, 0.3, 0.4, thr, gt)):.3f}  "
              f"fused fwd {t(lambda: head.map_update_loss(pred, tg, 0.3, 0.4, thr, gt, 5.0, 0.5, False)):.3f}  "
              f"fused fwd+bwd {t(lambda: head.map_update_loss(pred, tg, 0.3, 0.4, thr, gt, 5.0, 0.5, True)):.3f}", flush=True)
