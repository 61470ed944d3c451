"""yh_preprocess_rgb8 timing: CUDA events over repeated calls (image resident in HBM), the same through
YOLOInference.preprocess_image from a PIL image (host -> device copy inside), and the reference's torchvision
transform on the host.  Algorithmic bytes = input + 8-bit intermediate (written, read) + float32 output."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from PIL import Image
from torchvision import transforms
from yolo_v1_b200 import preprocess

tf = transforms.Compose([transforms.Resize((448, 448)), transforms.ToTensor(),
                         transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
rng = np.random.default_rng(0)
print(f"{'image':>12s} {'kernels us':>11s} {'GB/s':>8s} {'pil->gpu us':>12s} {'torchvision cpu us':>19s} threads={torch.get_num_threads()}")
for h, w in [(375, 500), (768, 1024), (2160, 3840), (6000, 8000)]:
    img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    d = torch.from_numpy(img).cuda()
    for _ in range(3):
        preprocess.preprocess_rgb8(d)
    torch.cuda.synchronize()
    reps = 50
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        preprocess.preprocess_rgb8(d)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    nbytes = h * w * 3 + (2 * h * 448 * 3 if w != 448 else 0) + 3 * 448 * 448 * 4
    pil = Image.fromarray(img)
    t0 = time.perf_counter()
    for _ in range(10):
        preprocess.preprocess_pil(pil, "cuda")
    torch.cuda.synchronize()
    us_pil = (time.perf_counter() - t0) * 1e5
    t0 = time.perf_counter()
    n_cpu = 10 if h * w < 4_000_000 else 3
    for _ in range(n_cpu):
        tf(pil)
    us_cpu = (time.perf_counter() - t0) * 1e6 / n_cpu
    print(f"{h:5d}x{w:<6d} {us:11.1f} {nbytes / us / 1e3:8.1f} {us_pil:12.1f} {us_cpu:19.1f}")

# a batch of equal-sized images: three launches for all of them
for n, h, w in [(64, 375, 500), (256, 375, 500)]:
    d = torch.from_numpy(rng.integers(0, 256, size=(n, h, w, 3), dtype=np.uint8)).cuda()
    for _ in range(3):
        preprocess.preprocess_rgb8(d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        preprocess.preprocess_rgb8(d)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    nbytes = n * (h * w * 3 + 2 * h * 448 * 3 + 3 * 448 * 448 * 4)
    print(f"batch {n} x {h}x{w}: {us:8.1f} us = {us / n:6.2f} us per image, {nbytes / us / 1e3:7.1f} GB/s")
