"""The head's last Linear fused with the loss (yh_head_linear_loss) against torch's Linear (cuBLAS) + this package's loss
kernel, CUDA events.

    python tools/time_head_linear.py [--rows 65536] [--dtype bf16|fp16|fp32] [--reps 10]

Prints one JSON line: ms and useful TFLOP/s (2 * M * 4096 * 1470) of the fused kernel (loss + narrow gradient, no
prediction written), of the plain layer (fp32 predictions written), of torch.nn.functional.linear with the layer's own
out_features = 1470 (row stride 2,940 B: not a multiple of 16) and with out_features padded to 1472, and of
linear + loss kernel."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import synth  # noqa: E402
from yolo_v1_b200.head_linear import linear_loss  # noqa: E402
from yolo_v1_b200.loss import loss_fwd_bwd  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=65536)
ap.add_argument("--dtype", default="bf16")
ap.add_argument("--reps", type=int, default=10)
args = ap.parse_args()
if args.dtype == "fp32":
    # the reference's default (non-AMP) step: fp32 operands; 3xTF32 (fp32's accuracy) and one TF32 pass against torch's fp32 Linear
    # with and without allow_tf32
    M = args.rows
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(M, 4096, device="cuda", generator=g) * 0.1
    w = torch.randn(1470, 4096, device="cuda", generator=g) * 0.02
    b = torch.rand(1470, device="cuda", generator=g) * 0.6
    tgt = synth.make_targets(M, seed=1, p_obj=0.055, size_mix=True).cuda()

    def timeit32(fn):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.reps

    def two():
        pred = torch.nn.functional.linear(x, w, b)
        return loss_fwd_bwd(pred.view(M, 7, 7, 30), tgt, 7, 2, 20, 5.0, 0.5, True)

    rows = {"fused_3xtf32": timeit32(lambda: linear_loss(x, w, b, tgt, precision="3xtf32")),
            "fused_tf32": timeit32(lambda: linear_loss(x, w, b, tgt, precision="tf32"))}
    torch.backends.cuda.matmul.allow_tf32 = False
    rows["torch_fp32_linear_plus_loss_kernel"] = timeit32(two)
    torch.backends.cuda.matmul.allow_tf32 = True
    rows["torch_allow_tf32_linear_plus_loss_kernel"] = timeit32(two)
    flops = 2.0 * M * 4096 * 1470
    print(json.dumps({"rows": M, "dtype": "fp32", "useful_flops": flops, "ms": {k: round(v, 4) for k, v in rows.items()},
                      "tflops_useful": {k: round(flops / v * 1e-9, 1) for k, v in rows.items()}}))
    sys.exit(0)
dt = {"bf16": torch.bfloat16, "fp16": torch.float16}[args.dtype]
M = args.rows


def timeit(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / args.reps


g = torch.Generator(device="cuda").manual_seed(0)
x = (torch.randn(M, 4096, device="cuda", generator=g) * 0.1).to(dt)
w = (torch.randn(1470, 4096, device="cuda", generator=g) * 0.02).to(dt)
b = torch.rand(1470, device="cuda", generator=g) * 0.6
w_pad = torch.zeros(1472, 4096, device="cuda", dtype=dt)
w_pad[:1470] = w
b_n, b_pad = b.to(dt), torch.zeros(1472, device="cuda", dtype=dt)
tgt = synth.make_targets(M, seed=1, p_obj=0.055, size_mix=True).cuda()
flops = 2.0 * M * 4096 * 1470


def two_calls():
    pred = torch.nn.functional.linear(x, w, b_n)
    return loss_fwd_bwd(pred.view(M, 7, 7, 30), tgt, 7, 2, 20, 5.0, 0.5, True)


rows = {
    "fused_linear_loss": timeit(lambda: linear_loss(x, w, b, tgt)),
    "fused_linear_loss_with_pred": timeit(lambda: linear_loss(x, w, b, tgt, want_pred=True)),
    "plain_linear_fp32_out": timeit(lambda: linear_loss(x, w, b, None)),
    "torch_linear_1470": timeit(lambda: torch.nn.functional.linear(x, w, b_n)),
    "torch_linear_padded_1472": timeit(lambda: torch.nn.functional.linear(x, w_pad, b_pad)),
    "torch_linear_1470_plus_loss_kernel": timeit(two_calls),
}
print(json.dumps({"rows": M, "dtype": args.dtype, "useful_flops": flops,
                  "ms": {k: round(v, 4) for k, v in rows.items()},
                  "tflops_useful": {k: round(flops / v * 1e-9, 1) for k, v in rows.items()}}))
