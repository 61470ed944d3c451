"""A few launches of the fused head Linear + loss for ncu: python tools/run_head_linear.py [--rows 65536] [--reps 3]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import synth  # noqa: E402
from yolo_v1_b200.head_linear import linear_loss  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=65536)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--plain", action="store_true")
args = ap.parse_args()
M = args.rows
g = torch.Generator(device="cuda").manual_seed(0)
x = (torch.randn(M, 4096, device="cuda", generator=g) * 0.1).to(torch.bfloat16)
w = (torch.randn(1470, 4096, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
b = torch.rand(1470, device="cuda", generator=g) * 0.6
tgt = synth.make_targets(M, seed=1, p_obj=0.055, size_mix=True).cuda()
for _ in range(args.reps):
    linear_loss(x, w, b, None if args.plain else tgt)
torch.cuda.synchronize()
