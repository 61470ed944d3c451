"""Generate golden vectors by running the UNMODIFIED reference (/root/reference/src).

Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py            # small fixtures (seconds..minutes)
    python tests/golden/make_golden.py --big      # + the 4,952-image compute() case (slow)

Outputs (committed): tests/golden/loss.npz, decode_nms.npz, map_small.npz, map_big.npz, encode.npz, validation.npz.
The reference is imported, never copied; only its outputs are stored.
"""
from __future__ import annotations

import argparse
import hashlib
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

S, B, C, K = 7, 2, 20, 98


def sha(*arrays) -> str:
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()[:16]


# ----------------------------------------------------------------------------- loss
def loss_case(pred: torch.Tensor, tgt: torch.Tensor, grad_scale: float = 1.0):
    crit = YOLOLoss(S=S, B=B, C=C)
    p = pred.clone().requires_grad_(True)
    total, d = crit(p, tgt)
    (total * grad_scale).backward()
    parts = np.array([d["total"], d["coord"], d["conf_obj"], d["conf_noobj"], d["class"]], np.float64)
    return parts, p.grad.detach()


def gen_loss(out):
    cases = {}
    for n in (1, 2, 16, 64):
        cases[f"uniform{n}"] = (synth.make_preds(n, "uniform", seed=100 + n), synth.make_targets(n, seed=200 + n))
    cases["noobj2"] = (synth.make_preds(2, "uniform", seed=11), torch.zeros(2, S, S, 30))
    cases["bothslots8"] = (synth.make_preds(8, "uniform", seed=12),
                           synth.make_targets(8, seed=13, p_obj=0.3, both_slots=0.5))
    # slot 1 filled but slot 0 empty: target box must come from slot 1 (loss.py:102-104)
    t = synth.make_targets(4, seed=14, p_obj=0.3)
    t2 = t.clone()
    t2[..., 5:10] = t[..., 0:5]
    t2[..., 0:5] = 0
    cases["slot1only4"] = (synth.make_preds(4, "uniform", seed=15), t2)
    # identical predicted boxes -> IoU tie -> box 0 responsible
    p = synth.make_preds(4, "uniform", seed=16)
    p[..., 5:9] = p[..., 0:4]
    cases["ioutie4"] = (p, synth.make_targets(4, seed=17, p_obj=0.3))
    cases["rawhead8"] = (synth.make_preds(8, "rawhead", seed=18), synth.make_targets(8, seed=19, p_obj=0.2))
    cases["ties8"] = (synth.make_preds(8, "ties", seed=20), synth.make_targets(8, seed=21, p_obj=0.2))
    tc = synth.make_targets(16, seed=22, p_obj=0.15, size_mix=True)
    cases["correlated16"] = (synth.make_preds(16, "correlated", seed=23, targets=tc), tc)
    # tiny widths below the 1e-6 clamp, and exactly at it
    p = synth.make_preds(4, "uniform", seed=24)
    p[..., 2] = 1e-7
    p[..., 8] = 1e-6
    cases["clamp4"] = (p, synth.make_targets(4, seed=25, p_obj=0.3))
    for name, (pred, tgt) in cases.items():
        parts, grad = loss_case(pred, tgt)
        out[f"{name}/pred"] = pred.numpy()
        out[f"{name}/tgt"] = tgt.numpy()
        out[f"{name}/parts"] = parts
        out[f"{name}/grad"] = grad.numpy()
    # GradScaler-style upstream gradient
    pred, tgt = cases["uniform16"]
    parts, grad = loss_case(pred, tgt, grad_scale=1024.0)
    out["uniform16_scaled/grad"] = grad.numpy()
    # half-precision predictions with fp32 targets (trainer.py:72 under autocast feeds these)
    for dt, nm in ((torch.float16, "f16"), (torch.bfloat16, "bf16")):
        pred, tgt = cases["uniform16"]
        ph = pred.to(dt)
        try:
            parts, grad = loss_case(ph, tgt)
            out[f"uniform16_{nm}/pred_as_f32"] = ph.float().numpy()
            out[f"uniform16_{nm}/parts"] = parts
            out[f"uniform16_{nm}/grad_as_f32"] = grad.float().numpy()
        except Exception as e:  # pragma: no cover
            print("half-precision reference run failed:", nm, e)


# ----------------------------------------------------------------------- decode + NMS
class _Shape(torch.nn.Module):
    """Stands in for the model: YOLOInference only reads .S and .B on this path."""

    def __init__(self):
        super().__init__()
        self.S, self.B = S, B


def metrics_decode_nms(metric: mAPMetric, pred: torch.Tensor):
    """Reference per-class NMS; returns ordered candidate indices + the kept tuples."""
    thr = metric.conf_threshold
    metric.conf_threshold = float("-inf")
    all98 = metric._parse_predictions(pred)
    metric.conf_threshold = thr
    assert len(all98) == K
    parsed = metric._parse_predictions(pred)
    mine = [d for d in all98 if d[1] > thr]
    assert parsed == mine, "candidate tagging does not reproduce the reference's filter"
    pos = {id(d): k for k, d in enumerate(all98)}
    kept = metric._apply_nms(mine)
    return [pos[id(d)] for d in kept], kept


def infer_decode_nms(inf: YOLOInference, pred: torch.Tensor, conf_thr: float, nms_thr: float):
    try:
        dets = inf.parse_predictions(pred, conf_thr)
    except ValueError:
        return None  # pydantic ValidationError: a kept box or confidence is outside [0,1]
    # tag by equality of (confidence, box) against the raw decode in candidate order
    m = mAPMetric(num_classes=C, conf_threshold=float("-inf"))
    all98 = m._parse_predictions(pred)
    ks = [k for k, d in enumerate(all98) if d[1] > conf_thr]
    assert len(ks) == len(dets)
    for k, d in zip(ks, dets):
        assert all98[k][1] == d.confidence and all98[k][0] == d.class_id
    pos = {id(d): k for k, d in zip(ks, dets)}
    kept = inf.non_max_suppression(dets, nms_thr)
    return [pos[id(d)] for d in kept]


def gen_decode_nms(out):
    inf = YOLOInference(_Shape(), device="cpu")
    n = 12
    for prof in ("uniform", "realistic", "ties", "rawhead", "correlated"):
        tgt = synth.make_targets(n, seed=31, p_obj=0.1, size_mix=True)
        pred = synth.make_preds(n, prof, seed=30, targets=tgt)
        out[f"{prof}/pred"] = pred.numpy()
        for conf_thr in (0.01, 0.3, 0.5):
            metric = mAPMetric(num_classes=C, conf_threshold=conf_thr, nms_threshold=0.4)
            mk = -np.ones((n, K), np.int32)
            mconf = np.zeros((n, K), np.float64)
            mcls = -np.ones((n, K), np.int32)
            mbox = np.zeros((n, K, 4), np.float64)
            ik = -np.ones((n, K), np.int32)
            raises = np.zeros(n, np.bool_)
            for i in range(n):
                ks, kept = metrics_decode_nms(metric, pred[i])
                mk[i, :len(ks)] = ks
                for q, d in enumerate(kept):
                    mcls[i, q], mconf[i, q], mbox[i, q] = d[0], d[1], d[2]
                r = infer_decode_nms(inf, pred[i], conf_thr, 0.4)
                if r is None:
                    raises[i] = True
                else:
                    ik[i, :len(r)] = r
            tag = f"{prof}/c{conf_thr}"
            out[f"{tag}/metrics_keep_k"] = mk
            out[f"{tag}/metrics_conf"] = mconf
            out[f"{tag}/metrics_cls"] = mcls
            out[f"{tag}/metrics_box"] = mbox
            out[f"{tag}/infer_keep_k"] = ik
            out[f"{tag}/infer_raises"] = raises


# ------------------------------------------------------------------------- mAP
class FlagTap:
    """Reads the tp arrays the reference builds inside _calculate_ap_for_class /
    _calculate_ap_for_size_class (metrics.py:388-413, 610-633) via sys.setprofile."""

    NAMES = ("_calculate_ap_for_class", "_calculate_ap_for_size_class")

    def __init__(self):
        self.calls = []  # (func, class_id, thr, tp array or None)

    def __call__(self, frame, event, arg):
        if event == "return" and frame.f_code.co_name in self.NAMES:
            loc = frame.f_locals
            tp = loc.get("tp")
            self.calls.append((frame.f_code.co_name, loc["class_id"], loc["iou_threshold"],
                               None if tp is None else np.array(tp, copy=True)))


def reference_map(pred, tgt, conf_thr, nms_thr=0.4, thresholds=None, want_flags=True):
    metric = mAPMetric(num_classes=C, iou_thresholds=thresholds, conf_threshold=conf_thr,
                       nms_threshold=nms_thr)
    keep_k = -np.ones((pred.shape[0], K), np.int32)
    for i in range(pred.shape[0]):
        ks, kept = metrics_decode_nms(metric, pred[i])
        keep_k[i, :len(ks)] = ks
        metric.all_predictions.append(kept)
        metric.all_ground_truths.append(metric._parse_ground_truth(tgt[i]))
    # sanity: same state as the reference's own update()
    chk = mAPMetric(num_classes=C, iou_thresholds=thresholds, conf_threshold=conf_thr, nms_threshold=nms_thr)
    chk.update(pred[:2], tgt[:2])
    assert chk.all_predictions == metric.all_predictions[:2]
    tap = FlagTap()
    sys.setprofile(tap)
    try:
        res = metric.compute()
    finally:
        sys.setprofile(None)
    thr = metric.iou_thresholds
    T = len(thr)
    flags = np.zeros((pred.shape[0], K), np.uint64)
    if want_flags:
        sizes = [s for s, nm in enumerate(("small", "medium", "large"), 1) if res[f"num_{nm}_objects"] > 0]
        subset_of_call = []
        seen_size = -1
        for func, c, t, tp in tap.calls:
            if func == "_calculate_ap_for_class":
                subset_of_call.append(0)
            else:
                if c == 0 and t == thr[0]:
                    seen_size += 1
                subset_of_call.append(sizes[seen_size])
        for (func, c, t, tp), s in zip(tap.calls, subset_of_call):
            if tp is None:
                continue
            ti = thr.index(t)
            # the reference's class_predictions order: (image, list position), stable sort by conf
            ident = [(i, q) for i, preds in enumerate(metric.all_predictions)
                     for q, d in enumerate(preds) if d[0] == c]
            conf = [metric.all_predictions[i][q][1] for i, q in ident]
            order = sorted(range(len(ident)), key=lambda r: conf[r], reverse=True)
            assert len(order) == len(tp)
            for r, o in enumerate(order):
                if tp[r] == 1:
                    i, q = ident[o]
                    flags[i, q] |= np.uint64(1) << np.uint64(s * T + ti)
    keys = sorted(res.keys())
    vals = np.array([float(res[k]) for k in keys], np.float64)
    return keep_k, flags, keys, vals


def gen_map_small(out):
    cases = {}
    tg = synth.make_targets(64, seed=41, p_obj=0.055, size_mix=True)
    cases["corr64"] = (synth.make_preds(64, "correlated", seed=40, targets=tg), tg, 0.01, None)
    tt = synth.make_targets(48, seed=43, p_obj=0.12)
    tt[..., 0:4] = torch.clamp(torch.round(tt[..., 0:4] * 8), min=1) / 8 * (tt[..., 4:5] > 0)
    cases["ties48"] = (synth.make_preds(48, "ties", seed=42), tt, 0.01, None)
    tu = synth.make_targets(32, seed=45, p_obj=0.1, size_mix=True)
    cases["uniform32"] = (synth.make_preds(32, "uniform", seed=44), tu, 0.3, None)
    tr = synth.make_targets(24, seed=47, p_obj=0.1, size_mix=True)
    cases["rawhead24"] = (synth.make_preds(24, "rawhead", seed=46), tr, 0.01, None)
    tc = synth.make_targets(40, seed=49, p_obj=0.08, size_mix=True)
    cases["corr40_thr3"] = (synth.make_preds(40, "correlated", seed=48, targets=tc), tc, 0.05, [0.3, 0.5, 0.9])
    # duplicate, tie-heavy detections competing for the same ground truths
    td = synth.make_targets(48, seed=53, p_obj=0.12)
    td[..., 0:4] = torch.clamp(torch.round(td[..., 0:4] * 8), min=1) / 8 * (td[..., 4:5] > 0)
    pd_ = synth.make_preds(48, "correlated", seed=52, targets=td)
    pd_[..., 5:10] = pd_[..., 0:5] + 0.04 * torch.randn(48, S, S, 5, generator=torch.Generator().manual_seed(54))
    pd_ = (torch.round(pd_ * 8) / 8).clamp(0.0, 1.0)
    cases["dupties48"] = (pd_, td, 0.01, None)
    for name, (pred, tgt, conf_thr, thrs) in cases.items():
        keep_k, flags, keys, vals = reference_map(pred, tgt, conf_thr, thresholds=thrs)
        out[f"{name}/pred"] = pred.numpy()
        out[f"{name}/tgt"] = tgt.numpy()
        out[f"{name}/conf_thr"] = np.float64(conf_thr)
        out[f"{name}/thresholds"] = np.array(thrs if thrs else [0.5 + 0.05 * i for i in range(10)], np.float64)
        out[f"{name}/keep_k"] = keep_k
        out[f"{name}/flags"] = flags
        out[f"{name}/keys"] = np.array(keys)
        out[f"{name}/vals"] = vals
        print(name, "mAP50:95 =", vals[keys.index("mAP50:95")], "kept/img =", (keep_k >= 0).sum() / pred.shape[0])


def gen_map_big(out, n, name):
    # synth.make_exact: integer hash + IEEE-exact operators only, so the GPU box regenerates these
    # inputs bit for bit whatever its CPU or thread count (the test FAILS on a hash mismatch)
    pred, tg = synth.make_exact(n, seed=50)
    keep_k, flags, keys, vals = reference_map(pred, tg, 0.01)
    out[f"{name}/n"] = np.int64(n)
    out[f"{name}/input_sha"] = np.array(sha(pred.numpy(), tg.numpy()))
    out[f"{name}/keep_k_sha"] = np.array(sha(keep_k))
    out[f"{name}/flags_sha"] = np.array(sha(flags))
    out[f"{name}/n_keep"] = (keep_k >= 0).sum(axis=1).astype(np.int32)
    out[f"{name}/keys"] = np.array(keys)
    out[f"{name}/vals"] = vals
    print(name, "mAP50:95 =", vals[keys.index("mAP50:95")])


def encode_cases():
    """(name, bboxes per sample, class ids per sample) for the target-encoder fixtures: VOC-like
    scenes, several boxes in one cell (first wins), boxes on the right/bottom edge (clamped to the
    last cell), slightly negative centres (Python's negative-index wrap), negative class ids."""
    rng = np.random.default_rng(20240611)
    cases = []
    for n_obj in (0, 1, 3, 8, 40):
        boxes = [[float(v) for v in rng.random(4)] for _ in range(n_obj)]
        cases.append((f"random{n_obj}", boxes, [int(c) for c in rng.integers(0, 20, n_obj)]))
    cases.append(("same_cell", [[0.50, 0.50, 0.2, 0.3], [0.52, 0.51, 0.1, 0.1], [0.55, 0.56, 0.4, 0.4]], [3, 4, 5]))
    cases.append(("edges", [[1.0, 1.0, 0.1, 0.1], [0.999999, 0.0, 0.2, 0.2], [0.0, 1.0, 0.3, 0.3], [1.2, 0.5, 0.1, 0.1]],
                  [0, 19, 7, 9]))
    cases.append(("negative", [[-0.05, 0.3, 0.1, 0.1], [0.3, -0.2, 0.2, 0.2], [-0.99, -0.99, 0.1, 0.1]], [1, 2, 3]))
    cases.append(("negative_class", [[0.2, 0.2, 0.1, 0.1], [0.7, 0.7, 0.1, 0.1]], [-1, -20]))
    cases.append(("thirds", [[1 / 3, 2 / 3, 0.1, 0.1], [1 / 7, 6 / 7, 0.2, 0.2], [3 / 7, 3 / 7, 0.3, 0.3]], [5, 6, 7]))
    return cases


def gen_encode(out):
    """VOCDetectionYOLO._encode_target (dataset.py:487-532) run unbound on a stand-in self."""
    from types import SimpleNamespace

    from yolo.dataset import VOCDetectionYOLO  # reference

    fake = SimpleNamespace(S=S, B=B, C=C)
    names = []
    for name, boxes, cls in encode_cases():
        t = VOCDetectionYOLO._encode_target(fake, boxes, cls)
        out[f"{name}/boxes"] = np.asarray(boxes, dtype=np.float64).reshape(-1, 4)
        out[f"{name}/cls"] = np.asarray(cls, dtype=np.int64)
        out[f"{name}/target"] = t.numpy()
        names.append(name)
    out["names"] = np.asarray(names)


def validation_cases():
    """Images for the pre-NMS validation of YOLOInference.predict (inference.py:133-136, 197-208)."""
    def blank():
        return torch.zeros(S, S, 30)

    def put(p, row, col, b, box, conf, cls):
        p[row, col, b * 5:b * 5 + 4] = torch.tensor(box)
        p[row, col, b * 5 + 4] = conf
        p[row, col, 10 + cls] = 1.0

    cases = {}
    # A (valid, conf 0.9) and B (same class, conf 0.8, centre x = (6 + 1.05) / 7 > 1) overlap with IoU 0.57:
    # NMS would drop B, but the reference validates B first and raises
    p = blank()
    put(p, 3, 6, 0, [0.90, 0.5, 0.30, 0.30], 0.9, 4)
    put(p, 3, 6, 1, [1.05, 0.5, 0.30, 0.30], 0.8, 4)
    cases["suppressed_x_gt_1"] = p
    # the same pair in range: B is suppressed, one detection comes back
    p = blank()
    put(p, 3, 6, 0, [0.90, 0.5, 0.30, 0.30], 0.9, 4)
    put(p, 3, 6, 1, [0.99, 0.5, 0.30, 0.30], 0.8, 4)
    cases["suppressed_in_range"] = p
    # two offenders: the EARLIER one in cell / box order raises (negative width at (0,2), conf > 1 at (5,1))
    p = blank()
    put(p, 5, 1, 0, [0.5, 0.5, 0.2, 0.2], 1.5, 7)
    put(p, 0, 2, 1, [0.5, 0.5, -0.1, 0.2], 0.7, 2)
    put(p, 0, 2, 0, [0.5, 0.5, 0.2, 0.2], 0.6, 2)
    cases["two_offenders"] = p
    # an offender below the confidence threshold is never built, so nothing raises
    p = blank()
    put(p, 2, 2, 0, [0.5, 0.5, 0.2, 0.2], 0.9, 1)
    put(p, 4, 4, 0, [0.5, 0.5, 1.7, 0.2], 0.4, 1)
    cases["offender_below_threshold"] = p
    # confidence product above 1 (class score 1.2 in a raw head)
    p = blank()
    put(p, 1, 1, 0, [0.5, 0.5, 0.2, 0.2], 0.95, 3)
    p[1, 1, 13] = 1.2
    cases["confidence_gt_1"] = p
    return cases


def gen_validation(out):
    """What the reference's parse_predictions + non_max_suppression do at conf 0.5 / nms 0.4."""
    import json

    inf = YOLOInference(_Shape(), device="cpu")
    names = []
    for name, pred in validation_cases().items():
        rec = {}
        try:
            dets = inf.parse_predictions(pred, 0.5)
            kept = inf.non_max_suppression(dets, 0.4)
            rec["raises"] = False
            rec["kept"] = [[d.class_id, d.confidence, d.bbox.x, d.bbox.y, d.bbox.width, d.bbox.height] for d in kept]
        except ValueError as e:   # pydantic.ValidationError
            err = e.errors()[0]
            rec["raises"] = True
            rec["title"] = e.title
            rec["type"] = err["type"]
            rec["loc"] = list(err["loc"])
            rec["input"] = float(err["input"])
        out[f"{name}/pred"] = pred.numpy()
        out[f"{name}/expect"] = np.array(json.dumps(rec))
        names.append(name)
        print(name, rec)
    out["names"] = np.asarray(names)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true")
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    warnings.simplefilter("ignore")
    torch.set_num_threads(1)
    jobs = {"loss": gen_loss, "decode_nms": gen_decode_nms, "map_small": gen_map_small, "encode": gen_encode,
            "validation": gen_validation}
    for name, fn in jobs.items():
        if a.only and a.only != name:
            continue
        out = {}
        fn(out)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
        print("wrote", name, len(out), "arrays")
    if a.big or a.only == "map_big":
        out = {}
        gen_map_big(out, 512, "exact512")
        gen_map_big(out, 4952, "exact4952")
        np.savez_compressed(os.path.join(HERE, "map_big.npz"), **out)
        print("wrote map_big")


if __name__ == "__main__":
    main()
