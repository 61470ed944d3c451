"""CPU-only checks: the C-ABI library loads and exports exactly what include/yolohead.h
declares; host-side logic of the drop-in classes that needs no kernel."""
import ctypes
import os
import re
import warnings

import numpy as np
import pytest
import torch

from yolo_v1_b200 import BoundingBox, Detection, YOLOLoss, _lib, mAPMetric
from yolo_v1_b200.metrics import _conf_from_key

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "yolohead.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(yh_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 15
    h = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(h, n), f"{n} declared in yolohead.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes table and header disagree"
    assert _lib.lib().yh_version() == 100


def test_workspace_queries_and_argument_errors_without_gpu():
    lib = _lib.lib()
    assert lib.yh_loss_workspace_bytes(64, 7, 2, 20) >= 256
    assert lib.yh_loss_workspace_bytes(64, 14, 2, 20) >= 256     # other shapes: the generic kernels
    assert lib.yh_loss_workspace_bytes(64, 7, 9, 20) == 0        # more than 8 boxes per cell: outside the build
    assert lib.yh_ap_workspace_bytes(1000, 20, 10) > 16 * 1000
    assert lib.yh_ap_workspace_bytes(1000, 20, 14) == 0          # too many thresholds
    assert lib.yh_compact_workspace_bytes(4952) >= 256
    # S = 9 is a supported (generic) shape now: the null pointers are what is wrong; S = 14 (392 candidates) is not
    rc = lib.yh_decode_nms(None, 1, 9, 2, 20, 0.1, 0.4, 0, 0, None, None, None, None, None, None, None, None)
    assert rc == -1
    rc = lib.yh_decode_nms(None, 1, 14, 2, 20, 0.1, 0.4, 0, 0, None, None, None, None, None, None, None, None)
    assert rc == -2 and b"S=14" in lib.yh_last_error()
    rc = lib.yh_loss_fwd_bwd(None, 0, None, 1, 7, 2, 20, 5.0, 0.5, None, None, None, 0, None)
    assert rc == -1


def test_cpu_tensors_fail_loudly():
    with pytest.raises(_lib.YoloHeadError):
        YOLOLoss()(torch.zeros(1, 7, 7, 30), torch.zeros(1, 7, 7, 30))
    with pytest.raises(_lib.YoloHeadError):
        mAPMetric(20).update(torch.zeros(1, 7, 7, 30), torch.zeros(1, 7, 7, 30))


def test_metric_constructor_and_reset():
    # mirrors the reference's tests/test_metrics.py:13-33
    m = mAPMetric(num_classes=20, iou_thresholds=[0.5])
    assert (m.num_classes, m.iou_thresholds, m.S, m.B) == (20, [0.5], 7, 2)
    assert mAPMetric(20).iou_thresholds == [0.5 + 0.05 * i for i in range(10)]
    assert mAPMetric(20, iou_thresholds=0.5).iou_thresholds == [0.5]
    assert mAPMetric(20).conf_threshold == 0.01 and mAPMetric(20).nms_threshold == 0.4
    m = mAPMetric(20)
    m.all_predictions.append([(0, 0.9, (0.5, 0.5, 0.2, 0.2))])
    m.all_ground_truths.append([(0, (0.5, 0.5, 0.2, 0.2))])
    m.reset()
    assert len(m.all_predictions) == 0 and len(m.all_ground_truths) == 0
    assert m.compute() == {"mAP50:95": 0.0, "mAP50": 0.0, "mAP75": 0.0, "precision": 0.0, "recall": 0.0}


def test_result_assembly_matches_reference_key_set():
    m = mAPMetric(20)
    ap = np.zeros((4, 20, 10))
    ap[0, 0, :] = 1.0
    gt = np.zeros((4, 20), np.int32)
    gt[0, 0] = 5
    gt[3, 0] = 5
    res = m._results(ap, 5, 5, gt, 5, 10, 20)
    assert res["AP50_class_0"] == 1.0 and res["AP50:95_class_0"] == pytest.approx(1.0)
    assert res["mAP50"] == pytest.approx(0.05) and res["mAP50:95"] == pytest.approx(0.05)
    assert res["precision"] == pytest.approx(1.0, abs=1e-5) and res["recall"] == pytest.approx(1.0, abs=1e-5)
    assert res["mAP50:95_small"] == 0.0 and res["num_large_objects"] == 5
    want = {"mAP50:95", "mAP50", "mAP75", "precision", "recall"}
    want |= {f"{p}_class_{c}" for p in ("AP50", "AP75", "AP50:95") for c in range(20)}
    want |= {f"{p}_{s}" for p in ("mAP50:95", "mAP50", "mAP75") for s in ("small", "medium", "large")}
    want |= {f"num_{s}_objects" for s in ("small", "medium", "large")}
    assert set(res) == want


def test_conf_key_roundtrip():
    conf = np.array([0.0, 1e-300, 0.3, 0.9999, 1.0, 7.5, np.inf, -0.25, -np.inf])
    bits = conf.view(np.uint64)
    orderable = np.where(bits >> np.uint64(63), ~bits, bits | np.uint64(1 << 63))
    key = ~orderable
    assert np.array_equal(_conf_from_key(key.view(np.int64)), conf)
    assert (np.diff(key[np.argsort(-conf)].astype(np.float64)) >= 0).all()   # key ascending = conf descending


def test_schemas_validate_like_the_reference():
    # mirrors the reference's tests/test_yolo.py:126-156
    b = BoundingBox(x=0.5, y=0.5, width=0.4, height=0.6)
    assert b.area == pytest.approx(0.24)
    assert b.to_corners() == pytest.approx((0.3, 0.2, 0.7, 0.8))
    assert b.to_pixel_coords(640, 480) == pytest.approx((192, 96, 448, 384), abs=1)
    assert str(b) == "(0.30, 0.20, 0.70, 0.80)"
    with pytest.raises(ValueError):
        BoundingBox(x=1.5, y=0.5, width=0.3, height=0.3)
    with pytest.raises(ValueError):
        BoundingBox(x=0.5, y=0.5, width=-0.1, height=0.3)
    with pytest.raises(ValueError):
        Detection(class_id=0, confidence=1.2, bbox=b)
    with pytest.raises(ValueError):
        Detection(class_id=-1, confidence=0.2, bbox=b)
    d = Detection(class_id=3, class_name="x", confidence=0.5, bbox=b)
    assert d == Detection(class_id=3, class_name="x", confidence=0.5, bbox=BoundingBox.from_corners(0.3, 0.2, 0.7, 0.8)) \
        or d.bbox.x == pytest.approx(0.5)


def test_loss_module_surface():
    crit = YOLOLoss(S=7, B=2, C=20, lambda_coord=5.0, lambda_noobj=0.5)
    assert (crit.S, crit.B, crit.C, crit.lambda_coord, crit.lambda_noobj) == (7, 2, 20, 5.0, 0.5)
    assert list(crit.parameters()) == [] and crit.state_dict() == {}
    a = torch.tensor([[[[[0.5, 0.5, 0.2, 0.2]]]]])
    assert YOLOLoss.compute_iou(a, a).item() == pytest.approx(1.0, abs=1e-4)
