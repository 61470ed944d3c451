"""Build container only (needs /root/reference): how much faster the CPU port that bench.py times on the GPU box
(oracle/yolo_oracle.c + the torch restatement of the loss) is than the UNMODIFIED reference, so that a reader can
chain reference -> port -> GPU.  Results are quoted in BASELINE.md section 3.

    python tools/port_vs_reference.py
"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root/reference/src")
sys.path.insert(0, ROOT)
from yolo.loss import YOLOLoss as RefLoss  # noqa: E402  (reference)
from yolo.metrics import mAPMetric as RefMetric  # noqa: E402  (reference)

from oracle import c_oracle, loss_oracle  # noqa: E402
from yolo_v1_b200 import synth  # noqa: E402

torch.set_num_threads(os.cpu_count() or 1)
print("host cores", os.cpu_count())
tgt = synth.make_targets(64, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(64, "uniform", seed=1234)
thr = [0.5 + 0.05 * i for i in range(10)]


def avg(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps


def ref_loss():
    p = pred.clone().requires_grad_(True)
    total, _ = RefLoss()(p, tgt)
    total.backward()


a, b = avg(ref_loss, 20), avg(lambda: loss_oracle.yolo_loss_and_grad(pred, tgt), 20)
print(f"loss fwd+bwd N=64: reference {a * 1e3:.2f} ms, port {b * 1e3:.2f} ms, ratio {a / b:.2f}")
for n in (128, 512):
    tg = synth.make_targets(n, seed=51, p_obj=0.055, size_mix=True)
    pr = synth.make_preds(n, "correlated", seed=50, targets=tg)
    m = RefMetric(num_classes=20, conf_threshold=0.01)
    t0 = time.perf_counter(); m.update(pr, tg); tu = time.perf_counter() - t0
    t0 = time.perf_counter(); m.compute(); tc = time.perf_counter() - t0
    row = [f"n={n}: reference update {tu / n * 1e3:.2f} ms/img, compute {tc:.2f} s"]
    for th in (1, os.cpu_count() or 1):
        t0 = time.perf_counter(); st = c_oracle.update_batch(pr.numpy(), tg.numpy(), 0.01, 0.4, n_threads=th); pu = time.perf_counter() - t0
        t0 = time.perf_counter(); c_oracle.compute(st, thr, n_threads=th); pc = time.perf_counter() - t0
        row.append(f"port {th} thread(s): update {pu / n * 1e6:.1f} us/img ({tu / pu:.0f}x), compute {pc * 1e3:.1f} ms ({tc / pc:.0f}x)")
    print(" | ".join(row))
