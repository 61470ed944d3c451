"""Launch yh_preprocess_rgb8 a few times on one large image (for ncu captures).

    ncu --set full --import-source on --clock-control none -k regex:pp_ -s 4 -c 4 -o gpurun_out/pp python tools/run_preprocess.py
"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import preprocess

h, w = int(os.environ.get("H", 6000)), int(os.environ.get("W", 8000))
img = torch.from_numpy(np.random.default_rng(0).integers(0, 256, size=(h, w, 3), dtype=np.uint8)).cuda()
for _ in range(3):
    out = preprocess.preprocess_rgb8(img)
torch.cuda.synchronize()
print("ok", tuple(out.shape))
