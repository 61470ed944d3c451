"""Randomised parity sweep: CUDA path vs the C oracle over many seeds, profiles, thresholds and batch splits.

    python tools/fuzz_parity.py [--rounds 200] [--seed0 1000]

Checks, per round: kept counts, kept candidate indices, TP/FP flag words (bit-exact), the result dict
(1e-12), the inference-order decode (bit-exact keep lists) and the loss parts / gradient (1e-5).
Not part of the test suite (needs minutes); the suites pin a fixed subset of the same comparisons.
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import c_oracle, loss_oracle, results  # noqa: E402
from yolo_v1_b200 import YOLOLoss, head, mAPMetric, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rounds", type=int, default=200)
ap.add_argument("--seed0", type=int, default=1000)
a = ap.parse_args()
dev = torch.device("cuda")
rng = np.random.default_rng(a.seed0)
profiles = ["uniform", "realistic", "correlated", "ties", "rawhead"]
bad = 0
for r in range(a.rounds):
    seed = a.seed0 + r
    n = int(rng.integers(1, 400))
    profile = profiles[r % len(profiles)]
    conf = float(rng.choice([0.0, 0.01, 0.05, 0.3, 0.5, 0.9]))
    nms = float(rng.choice([0.0, 0.1, 0.4, 0.5, 0.75, 1.0]))
    T = int(rng.integers(1, 14))
    thr = sorted(set(float(x) for x in np.round(rng.uniform(0.05, 0.95, T), 2))) if r % 3 else \
        [0.5 + 0.05 * i for i in range(10)]
    p_obj = float(rng.choice([0.0, 0.02, 0.055, 0.3, 1.0]))
    tgt = synth.make_targets(n, seed=seed, p_obj=p_obj, size_mix=bool(r % 2), both_slots=0.1 * (r % 2))
    pred = synth.make_preds(n, profile, seed=seed + 7, targets=tgt)
    tag = f"round {r}: n={n} {profile} conf={conf} nms={nms} T={len(thr)} p_obj={p_obj}"
    try:
        st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), conf, nms, n_threads=8)
        ref = c_oracle.compute(st, thr, n_threads=8)
        m = mAPMetric(num_classes=20, iou_thresholds=thr, conf_threshold=conf, nms_threshold=nms)
        cut = int(rng.integers(1, n + 1))
        for lo, hi in ((0, cut), (cut, n)):
            if hi > lo:
                m.update(pred[lo:hi].to(dev), tgt[lo:hi].to(dev))
        n_keep = torch.cat([c.n_keep for c in m._chunks]).cpu().numpy()
        assert (n_keep == st["n_det"]).all(), "kept counts"
        ks = torch.cat([c.rec_k for c in m._chunks]).cpu().numpy()
        fl = torch.cat([c.rec_flags for c in m._chunks]).cpu().numpy().view(np.uint64) & np.uint64((1 << 56) - 1)
        for i in range(n):
            mine = {int(ks[i, q]): int(fl[i, q]) for q in range(n_keep[i])}
            want = {int(st["det_k"][i, q]): int(ref["flags"][i, q]) for q in range(st["n_det"][i])}
            assert mine == want, f"flags of image {i}"
        res = m.compute()
        want = results.assemble(ref["ap"], ref["overall"], ref["sizes"], thr, 20, n)
        assert sorted(res) == sorted(want), "key set"
        for k in want:
            assert abs(float(res[k]) - float(want[k])) <= 1e-12 * max(1.0, abs(float(want[k]))), k
        if profile != "rawhead":
            d = head.decode_nms(pred.to(dev), conf, nms, head.IOU_INFER, head.ORDER_GLOBAL_CONF)
            o = c_oracle.update_batch(pred.numpy(), None, conf, nms, iou_mode=c_oracle.IOU_INFER,
                                      order=c_oracle.ORDER_GLOBAL_CONF, n_threads=8)
            assert (d.n_keep.cpu().numpy() == o["n_det"]).all(), "inference kept counts"
            kk = d.keep_k.cpu().numpy()
            for i in range(n):
                assert kk[i, :o["n_det"][i]].tolist() == o["det_k"][i, :o["n_det"][i]].tolist(), f"inference order {i}"
        parts_ref, grad_ref = loss_oracle.yolo_loss_and_grad(pred, tgt)
        p = pred.to(dev).requires_grad_(True)
        total, dct = YOLOLoss()(p, tgt.to(dev))
        total.backward()
        got = np.array([dct[k] for k in ("total", "coord", "conf_obj", "conf_noobj", "class")])
        assert np.allclose(got, parts_ref.double().numpy(), rtol=1e-5, atol=1e-7), "loss parts"
        g, gr = p.grad.cpu().numpy(), grad_ref.numpy()
        scale = max(1e-30, float(np.abs(gr).max()))
        assert np.abs(g - gr).max() <= 2e-5 * scale + 1e-5 * float(np.abs(gr).max()), "loss gradient"
    except AssertionError as e:
        bad += 1
        print("MISMATCH", tag, "->", e)
    except Exception as e:  # noqa: BLE001
        bad += 1
        print("ERROR", tag, "->", type(e).__name__, e)
print(f"{a.rounds} rounds, {bad} failures")
sys.exit(1 if bad else 0)
