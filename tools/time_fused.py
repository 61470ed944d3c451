"""The fused loss + update kernel against the two separate kernels (CUDA events, 65,536 images)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import head, synth  # noqa: E402
from yolo_v1_b200.loss import loss_fwd_bwd  # noqa: E402

thr = [0.5 + 0.05 * i for i in range(10)]


def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for prof, conf in (("uniform", 0.3), ("realistic", 0.3), ("correlated", 0.01)):
    tgt = synth.make_targets(65536, seed=1234, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(65536, prof, seed=1234, targets=tgt).cuda()
    tgt = tgt.cuda()
    gt = torch.zeros(80, dtype=torch.int32, device="cuda")
    a = t(lambda: loss_fwd_bwd(pred, tgt, 7, 2, 20, 5.0, 0.5, True))
    b = t(lambda: head.map_update(pred, tgt, conf, 0.4, thr, gt))
    c = t(lambda: head.map_update_loss(pred, tgt, conf, 0.4, thr, gt, 5.0, 0.5, True))
    d = t(lambda: head.map_update_loss(pred, tgt, conf, 0.4, thr, gt, 5.0, 0.5, False))
    print(f"{prof}: loss {a:.3f} + map_update {b:.3f} = {a + b:.3f} ms; fused fwd+bwd {c:.3f} ms, fused fwd only {d:.3f} ms", flush=True)
