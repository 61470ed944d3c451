"""Half-precision predictions: deviation of the loss kernel from the reference, per loss term.

The reference under autocast (trainer.py:72) feeds fp16 / bf16 predictions and fp32 targets to YOLOLoss; type
promotion makes every op that touches a target fp32, but the ops that touch ONLY predictions stay narrow and round
there: `boxes1[..., 2] / 2`, `x -/+ w/2` (the corners of the predicted boxes), `w * h` (their areas) in
compute_iou (loss.py:191-205), `torch.clamp(pred_wh, min=1e-6)` and `torch.sqrt` of it (loss.py:131-132), and
`pred_conf ** 2`'s operand in the no-object term (loss.py:150-153).  The kernel widens the staged values to fp32 once
and computes everything in fp32, so it is CLOSER to the exact value than the reference is; the difference per term is
tabulated here against tests/golden/loss.npz (reference outputs for the same narrow inputs)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import YOLOLoss  # noqa: E402

g = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "loss.npz"))
tgt = torch.from_numpy(g["uniform16/tgt"]).cuda()
names = ("total", "coord", "conf_obj", "conf_noobj", "class")
print("| predictions | term | reference (narrow ops) | kernel (fp32 inside) | relative deviation |\n|---|---|---|---|---|")
for dt, nm in ((torch.float32, None), (torch.float16, "f16"), (torch.bfloat16, "bf16")):
    if nm is None:
        pred = torch.from_numpy(g["uniform16/pred"]).cuda()
        ref_parts, ref_grad = g["uniform16/parts"], g["uniform16/grad"]
    else:
        pred = torch.from_numpy(g[f"uniform16_{nm}/pred_as_f32"]).cuda().to(dt)
        ref_parts, ref_grad = g[f"uniform16_{nm}/parts"], g[f"uniform16_{nm}/grad_as_f32"]
    p = pred.clone().requires_grad_(True)
    total, parts = YOLOLoss()(p, tgt)
    total.backward()
    for i, k in enumerate(names):
        print(f"| {str(dt).split('.')[-1]} | {k} | {ref_parts[i]:.7g} | {parts[k]:.7g} | {abs(parts[k] - ref_parts[i]) / max(abs(ref_parts[i]), 1e-30):.2e} |")
    gk = p.grad.float().cpu().numpy()
    den = np.abs(ref_grad).sum()
    print(f"| {str(dt).split('.')[-1]} | gradient | sum abs {den:.6g} | | sum abs diff / sum abs = {np.abs(gk - ref_grad).sum() / den:.2e}, max abs diff {np.abs(gk - ref_grad).max():.2e} |")
