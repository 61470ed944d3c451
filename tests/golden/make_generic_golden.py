"""Golden vectors for head shapes other than S=7, B=2, C=20, from the UNMODIFIED reference (/root/reference/src).

    python tests/golden/make_generic_golden.py        ->  tests/golden/generic.npz

The reference's classes take S, B and the class count as constructor arguments (loss.py:28-53,
metrics.py:43-71, inference.py:170-171); these fixtures hold, per shape: the loss parts + gradient, the
ordered keep lists of both NMS flavours, and the compute() dict.  Build container only.
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/src")

from yolo.inference import YOLOInference  # noqa: E402  (reference)
from yolo.loss import YOLOLoss  # noqa: E402  (reference)
from yolo.metrics import mAPMetric  # noqa: E402  (reference)

from yolo_v1_b200 import synth  # noqa: E402

SHAPES = [(7, 2, 3), (7, 2, 1), (7, 2, 32), (4, 3, 5), (11, 2, 20), (3, 1, 2), (5, 8, 7), (7, 1, 20)]


class _Shape(torch.nn.Module):
    def __init__(self, S, B):
        super().__init__()
        self.S, self.B = S, B


def tag_kept(all_dets, kept):
    pos = {id(d): k for k, d in enumerate(all_dets)}
    return [pos[id(d)] for d in kept]


def main():
    warnings.simplefilter("ignore")
    torch.set_num_threads(1)
    out = {}
    names = []
    for S, B, C in SHAPES:
        name = f"S{S}B{B}C{C}"
        names.append(name)
        K = S * S * B
        n = 10
        seed = 1000 * S + 10 * B + C
        tgt = synth.make_targets(n, seed=seed, S=S, B=B, C=C, p_obj=0.25, size_mix=True, both_slots=0.3 if B > 1 else 0.0)
        for prof in ("uniform", "ties"):
            pred = synth.make_preds(n, prof, seed=seed + 1, S=S, B=B, C=C)
            if prof == "uniform":
                # some true positives: copy target boxes into slot 0 of object cells, sharpen the class
                obj = tgt[..., 4] > 0
                pred[..., 0:4] = torch.where(obj[..., None], (tgt[..., 0:4] + 0.03 * torch.randn(n, S, S, 4, generator=torch.Generator().manual_seed(seed))).clamp(0, 1), pred[..., 0:4])
                pred[..., B * 5:] = torch.where(obj[..., None], 0.2 * pred[..., B * 5:] + 0.8 * tgt[..., B * 5:], pred[..., B * 5:])
            tag = f"{name}/{prof}"
            out[f"{tag}/pred"] = pred.numpy()
            out[f"{tag}/tgt"] = tgt.numpy()
            # ---- loss
            crit = YOLOLoss(S=S, B=B, C=C)
            p = pred.clone().requires_grad_(True)
            total, d = crit(p, tgt)
            total.backward()
            out[f"{tag}/loss_parts"] = np.array([d["total"], d["coord"], d["conf_obj"], d["conf_noobj"], d["class"]], np.float64)
            out[f"{tag}/loss_grad"] = p.grad.numpy()
            # ---- metrics flavour: ordered keep lists + compute()
            conf_thr = 0.2
            metric = mAPMetric(num_classes=C, conf_threshold=conf_thr, nms_threshold=0.4, S=S, B=B)
            mk = -np.ones((n, K), np.int32)
            for i in range(n):
                metric.conf_threshold = float("-inf")
                all_k = metric._parse_predictions(pred[i])
                metric.conf_threshold = conf_thr
                assert len(all_k) == K
                mine = [dd for dd in all_k if dd[1] > conf_thr]
                assert mine == metric._parse_predictions(pred[i])
                kept = metric._apply_nms(mine)
                ks = tag_kept(all_k, kept)
                mk[i, :len(ks)] = ks
            out[f"{tag}/metrics_keep_k"] = mk
            m2 = mAPMetric(num_classes=C, conf_threshold=conf_thr, nms_threshold=0.4, S=S, B=B)
            m2.update(pred, tgt)
            res = m2.compute()
            keys = sorted(res.keys())
            out[f"{tag}/keys"] = np.array(keys)
            out[f"{tag}/vals"] = np.array([float(res[k]) for k in keys], np.float64)
            out[f"{tag}/n_gt"] = np.array([len(g) for g in m2.all_ground_truths], np.int32)
            # ---- inference flavour
            inf = YOLOInference(_Shape(S, B), device="cpu")
            ik = -np.ones((n, K), np.int32)
            raises = np.zeros(n, np.bool_)
            for i in range(n):
                try:
                    dets = inf.parse_predictions(pred[i], conf_thr)
                except ValueError:
                    raises[i] = True
                    continue
                mall = mAPMetric(num_classes=C, conf_threshold=float("-inf"), S=S, B=B)._parse_predictions(pred[i])
                ks = [k for k, dd in enumerate(mall) if dd[1] > conf_thr]
                assert len(ks) == len(dets)
                pos = {id(dd): k for k, dd in zip(ks, dets)}
                kept = inf.non_max_suppression(dets, 0.4)
                r = [pos[id(dd)] for dd in kept]
                ik[i, :len(r)] = r
            out[f"{tag}/infer_keep_k"] = ik
            out[f"{tag}/infer_raises"] = raises
            print(tag, "loss", d["total"], "mAP50", res.get("mAP50"), "kept/img", (mk >= 0).sum() / n, "raises", int(raises.sum()))
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "generic.npz"), **out)
    print("wrote generic.npz")


if __name__ == "__main__":
    main()
