"""A few launches of the fused head Linear + loss for ncu:
python tools/run_head_linear.py [--rows 65536] [--reps 3] [--dtype bf16|fp16|fp32|tf32] [--plain]"""
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
ap.add_argument("--dtype", default="bf16")
ap.add_argument("--plain", action="store_true")
args = ap.parse_args()
M = args.rows
dt = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32, "tf32": torch.float32}[args.dtype]
g = torch.Generator(device="cuda").manual_seed(0)
x = (torch.randn(M, 4096, device="cuda", generator=g) * 0.1).to(dt)
w = (torch.randn(1470, 4096, device="cuda", generator=g) * 0.02).to(dt)
b = torch.rand(1470, device="cuda", generator=g) * 0.6
tgt = synth.make_targets(M, seed=1, p_obj=0.055, size_mix=True).cuda()
for _ in range(args.reps):
    linear_loss(x, w, b, None if args.plain else tgt, precision="tf32" if args.dtype == "tf32" else "3xtf32")
torch.cuda.synchronize()
