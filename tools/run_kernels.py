"""Launch each hot kernel a few times on the bench workload (for ncu captures).

    ncu --set full --import-source on --clock-control none -k regex:map_update -s 2 -c 1 \
        -o gpurun_out/map_update python tools/run_kernels.py
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import head, synth  # noqa: E402
from yolo_v1_b200.loss import loss_fwd_bwd  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--images", type=int, default=65536)
ap.add_argument("--profile", default="uniform")
ap.add_argument("--conf", type=float, default=0.3)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
tgt = synth.make_targets(a.images, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(a.images, a.profile, seed=1234, targets=tgt).cuda()
tgt = tgt.cuda()
thr = [0.5 + 0.05 * i for i in range(10)]
for _ in range(a.reps):
    loss_fwd_bwd(pred, tgt, 7, 2, 20, 5.0, 0.5, True)
    gt = torch.zeros(80, dtype=torch.int32, device="cuda")
    ch = head.map_update(pred, tgt, a.conf, 0.4, thr, gt)
    head.decode_nms(pred, a.conf, 0.4, head.IOU_INFER, head.ORDER_GLOBAL_CONF, want_boxes=False)
    k, f, off = head.compact_records(ch)
    m = int(off[-1].item())
    head.ap_reduce(k[:m].clone(), f[:m].clone(), m, gt, 20, 10)
torch.cuda.synchronize()
print("records", m)
