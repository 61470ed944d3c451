"""Assemble the reference's result dict from the oracle's AP table.

TEST INFRASTRUCTURE ONLY.  Mirrors /root/reference/src/yolo/metrics.py:117-171
(keys, np.mean call sites and their element order) and :531-564 (size keys).
"""
from __future__ import annotations

import numpy as np


def assemble(ap, overall, sizes, thresholds, num_classes, n_images):
    if n_images == 0:  # metrics.py:117-124
        return {"mAP50:95": 0.0, "mAP50": 0.0, "mAP75": 0.0, "precision": 0.0, "recall": 0.0}
    thr = list(thresholds)
    res = {}
    per_thr = {t: [] for t in thr}
    for c in range(num_classes):
        row = []
        for ti, t in enumerate(thr):
            a = float(ap[0, c, ti])
            per_thr[t].append(a)
            row.append(a)
            if t == 0.5:
                res[f"AP50_class_{c}"] = a
            elif t == 0.75:
                res[f"AP75_class_{c}"] = a
        res[f"AP50:95_class_{c}"] = np.mean(row)
    if 0.5 in thr:
        res["mAP50"] = np.mean(per_thr[0.5])
    if 0.75 in thr:
        res["mAP75"] = np.mean(per_thr[0.75])
    res["mAP50:95"] = np.mean([a for aps in per_thr.values() for a in aps])
    tp, fp, gt = (int(v) for v in overall)
    res["precision"] = tp / (tp + fp + 1e-6)
    res["recall"] = tp / (gt + 1e-6)
    for s, name in enumerate(("small", "medium", "large"), start=1):
        if int(sizes[s - 1]) == 0:
            res[f"mAP50:95_{name}"] = 0.0
            res[f"mAP50_{name}"] = 0.0
            res[f"mAP75_{name}"] = 0.0
            continue
        per = {t: [] for t in thr}
        for c in range(num_classes):
            for ti, t in enumerate(thr):
                per[t].append(float(ap[s, c, ti]))
        if 0.5 in thr:
            res[f"mAP50_{name}"] = np.mean(per[0.5])
        if 0.75 in thr:
            res[f"mAP75_{name}"] = np.mean(per[0.75])
        res[f"mAP50:95_{name}"] = np.mean([a for aps in per.values() for a in aps])
    res["num_small_objects"] = int(sizes[0])
    res["num_medium_objects"] = int(sizes[1])
    res["num_large_objects"] = int(sizes[2])
    return res
