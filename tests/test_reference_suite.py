"""The scenarios of the reference's own hot-path tests, run against the drop-in.

Each test names the reference test it restates (tests/test_metrics.py and
tests/test_yolo.py of mattiaskvist/yolo-v1, file:line in the docstring).  Like the
reference's tests, these hand HOST tensors to the metric and to the inference
engine -- the drop-in copies them to the GPU -- so the unmodified reference suite
runs against `yolo_v1_b200` on any machine with a CUDA device.
"""
import warnings

import pytest
import torch
import torch.nn as nn

from yolo_v1_b200 import BoundingBox, Detection, YOLOInference, mAPMetric

pytestmark = pytest.mark.gpu


def _cell(t, img, row, col, box, conf, cls, prob=1.0):
    t[img, row, col, 0:4] = torch.tensor(box)
    t[img, row, col, 4] = conf
    t[img, row, col, 10 + cls] = prob


# ------------------------------------------------------------------ tests/test_metrics.py
def test_metric_attributes_after_construction():
    """test_initialization, test_metrics.py:13-19"""
    m = mAPMetric(num_classes=20, iou_thresholds=[0.5])
    assert (m.num_classes, m.iou_thresholds, m.S, m.B) == (20, [0.5], 7, 2)


def test_reset_empties_the_public_lists():
    """test_reset, test_metrics.py:21-33"""
    m = mAPMetric(num_classes=20)
    m.all_predictions.append([(0, 0.9, (0.5, 0.5, 0.2, 0.2))])
    m.all_ground_truths.append([(0, (0.5, 0.5, 0.2, 0.2))])
    m.reset()
    assert len(m.all_predictions) == 0 and len(m.all_ground_truths) == 0


def test_metric_iou_identical_disjoint_partial_and_degenerate():
    """test_iou_calculation, test_metrics.py:35-55; test_iou_edge_cases, :208-222"""
    m = mAPMetric(num_classes=20)
    assert m._calculate_iou((0.5, 0.5, 0.2, 0.2), (0.5, 0.5, 0.2, 0.2)) == pytest.approx(1.0, abs=1e-5)
    assert m._calculate_iou((0.2, 0.2, 0.1, 0.1), (0.8, 0.8, 0.1, 0.1)) == 0.0
    assert 0 < m._calculate_iou((0.5, 0.5, 0.4, 0.4), (0.6, 0.6, 0.4, 0.4)) < 1
    assert m._calculate_iou((0.5, 0.5, 0.0, 0.0), (0.5, 0.5, 0.2, 0.2)) == 0.0
    assert m._calculate_iou((0.5, 0.5, 0.0, 0.0), (0.5, 0.5, 0.0, 0.0)) == 0.0


def test_parse_predictions_finds_the_confident_box():
    """test_parse_predictions, test_metrics.py:57-76"""
    m = mAPMetric(num_classes=20, conf_threshold=0.1, S=7, B=2)
    pred = torch.zeros(1, 7, 7, 30)
    _cell(pred, 0, 3, 3, [0.5, 0.5, 0.3, 0.3], 0.9, 0)
    dets = m._parse_predictions(pred[0])
    assert len(dets) > 0
    cls, conf, box = dets[0]
    assert cls == 0 and conf > 0.1 and len(box) == 4


def test_parse_ground_truth_finds_the_object_cell():
    """test_parse_ground_truth, test_metrics.py:78-96"""
    m = mAPMetric(num_classes=20, S=7, B=2)
    tgt = torch.zeros(1, 7, 7, 30)
    _cell(tgt, 0, 3, 3, [0.5, 0.5, 0.3, 0.3], 1.0, 0)
    gts = m._parse_ground_truth(tgt[0])
    assert len(gts) == 1 and gts[0][0] == 0


def test_metric_nms_three_to_two():
    """test_nms, test_metrics.py:98-117"""
    m = mAPMetric(num_classes=20, nms_threshold=0.5)
    kept = m._apply_nms([(0, 0.9, (0.5, 0.5, 0.2, 0.2)), (0, 0.8, (0.52, 0.52, 0.2, 0.2)),
                         (1, 0.85, (0.7, 0.7, 0.15, 0.15))])
    assert len(kept) == 2
    zero = [d for d in kept if d[0] == 0]
    assert len(zero) == 1 and zero[0][1] == 0.9


def test_update_compute_key_set_and_ranges():
    """test_update_and_compute, test_metrics.py:119-156"""
    m = mAPMetric(num_classes=20, S=7, B=2)
    pred, tgt = torch.zeros(2, 7, 7, 30), torch.zeros(2, 7, 7, 30)
    for i in range(2):
        _cell(pred, i, 3, 3, [0.5, 0.5, 0.3, 0.3], 0.8, 0)
        _cell(tgt, i, 3, 3, [0.5, 0.5, 0.3, 0.3], 1.0, 0)
    m.update(pred, tgt)
    res = m.compute()
    for k in ("mAP50:95", "mAP50", "mAP75", "precision", "recall"):
        assert k in res
    for c in range(20):
        assert f"AP50_class_{c}" in res and f"AP75_class_{c}" in res
        assert 0.0 <= res[f"AP50:95_class_{c}"] <= 1.0


def test_perfect_predictions_score_one():
    """test_perfect_predictions, test_metrics.py:158-183"""
    m = mAPMetric(num_classes=20, S=7, B=2, iou_thresholds=[0.5])
    pred, tgt = torch.zeros(5, 7, 7, 30), torch.zeros(5, 7, 7, 30)
    for i in range(5):
        _cell(pred, i, 3, 3, [0.5, 0.5, 0.3, 0.3], 1.0, 0)
        _cell(tgt, i, 3, 3, [0.5, 0.5, 0.3, 0.3], 1.0, 0)
    m.update(pred, tgt)
    res = m.compute()
    for k in ("AP50_class_0", "AP50:95_class_0", "precision", "recall"):
        assert res[k] == pytest.approx(1.0, abs=1e-5), k


def test_nothing_above_threshold_gives_zero_recall():
    """test_no_predictions, test_metrics.py:185-205"""
    m = mAPMetric(num_classes=20, conf_threshold=0.9)
    pred, tgt = torch.zeros(2, 7, 7, 30), torch.zeros(2, 7, 7, 30)
    pred[:, 3, 3, 4] = 0.1
    for i in range(2):
        _cell(tgt, i, 3, 3, [0.5, 0.5, 0.3, 0.3], 1.0, 0)
    m.update(pred, tgt)
    assert m.compute()["recall"] == 0.0


# ------------------------------------------------------------------ tests/test_yolo.py
class _Head(nn.Module):
    """Stand-in for YOLOv1: the inference engine only reads .S and .B and calls forward."""

    def __init__(self):
        super().__init__()
        self.S, self.B, self.C = 7, 2, 20
        self.out = nn.Parameter(torch.zeros(1, 7, 7, 30))

    def forward(self, x):
        return self.out.expand(x.shape[0], -1, -1, -1)


@pytest.fixture(params=["cpu", "cuda"])
def engine(request):
    """device="cpu" is what the reference's fixture passes (test_yolo.py:24-27): the model
    then runs on the host and only the post-processing goes to the GPU."""
    return YOLOInference(_Head(), device=request.param)


def _det(cls, conf, x, y, w, h):
    return Detection(class_id=cls, class_name=f"class_{cls}", confidence=conf,
                     bbox=BoundingBox(x=x, y=y, width=w, height=h))


def test_engine_construction(engine):
    """test_initialization, test_yolo.py:49-56"""
    assert engine.device in ("cpu", "cuda")
    assert hasattr(engine, "transform") and not engine.model.training


def test_predict_on_an_image_file_and_on_a_missing_one(engine, tmp_path):
    """test_predict_returns_list / test_predict_with_thresholds / test_predict_invalid_image_path, test_yolo.py:58-83"""
    from PIL import Image

    path = tmp_path / "red.jpg"
    Image.new("RGB", (448, 448), color="red").save(path)
    dets = engine.predict(str(path))
    assert isinstance(dets, list) and all(isinstance(d, Detection) for d in dets)
    lo = engine.predict(str(path), conf_threshold=0.1, nms_threshold=0.3)
    hi = engine.predict(str(path), conf_threshold=0.9, nms_threshold=0.3)
    assert len(hi) <= len(lo)
    with pytest.raises(FileNotFoundError):
        engine.predict("nonexistent_image.jpg")


def test_parse_predictions_returns_validated_detections(engine):
    """test_parse_predictions_shape, test_yolo.py:85-109"""
    pred = torch.zeros(7, 7, 30)
    pred[2, 3, 0:5] = torch.tensor([0.5, 0.5, 0.3, 0.3, 0.9])
    pred[2, 3, 10] = 0.8
    dets = engine.parse_predictions(pred, conf_threshold=0.5)
    assert isinstance(dets, list) and len(dets) == 1
    d = dets[0]
    assert isinstance(d.class_id, int) and isinstance(d.confidence, float) and isinstance(d.bbox, BoundingBox)
    assert 0 <= d.bbox.x <= 1 and 0 <= d.bbox.y <= 1 and 0 <= d.bbox.width <= 1 and 0 <= d.bbox.height <= 1
    assert d.confidence == pytest.approx(0.9 * 0.8, rel=1e-6)


def test_confidence_threshold_filters(engine):
    """test_parse_predictions_confidence_threshold, test_yolo.py:111-124"""
    pred = torch.zeros(7, 7, 30)
    pred[0, 0, 0:5] = torch.tensor([0.5, 0.5, 0.3, 0.3, 0.9])
    pred[0, 0, 10] = 0.8
    assert len(engine.parse_predictions(pred, conf_threshold=0.5)) > 0
    assert len(engine.parse_predictions(pred, conf_threshold=0.9)) == 0


def test_out_of_range_boxes_raise_like_the_schema(engine):
    """test_detection_bbox_validation, test_yolo.py:126-138; the same validation fires inside parse_predictions"""
    assert BoundingBox(x=0.5, y=0.5, width=0.3, height=0.3).area == pytest.approx(0.09, abs=1e-6)
    with pytest.raises(ValueError):
        BoundingBox(x=1.5, y=0.5, width=0.3, height=0.3)
    with pytest.raises(ValueError):
        BoundingBox(x=0.5, y=0.5, width=-0.1, height=0.3)
    pred = torch.zeros(7, 7, 30)
    pred[1, 1, 0:5] = torch.tensor([0.5, 0.5, 1.7, 0.3, 0.9])   # width > 1 on a confident box
    pred[1, 1, 10] = 0.9
    with pytest.raises(ValueError):
        engine.parse_predictions(pred, conf_threshold=0.5)


def test_corner_and_pixel_conversion():
    """test_detection_coordinate_conversion, test_yolo.py:140-156"""
    b = BoundingBox(x=0.5, y=0.5, width=0.4, height=0.6)
    assert b.to_corners() == pytest.approx((0.3, 0.2, 0.7, 0.8), abs=1e-6)
    assert b.to_pixel_coords(640, 480) == pytest.approx((192, 96, 448, 384), abs=1)


def test_engine_iou(engine):
    """test_iou_identical_boxes / no_overlap / partial_overlap / symmetry, test_yolo.py:196-227"""
    a = BoundingBox(x=0.5, y=0.5, width=0.2, height=0.2)
    assert engine.iou(a, a) == pytest.approx(1.0, abs=1e-3)
    assert engine.iou(BoundingBox(x=0.2, y=0.2, width=0.1, height=0.1),
                      BoundingBox(x=0.8, y=0.8, width=0.1, height=0.1)) == pytest.approx(0.0, abs=1e-6)
    p, q = BoundingBox(x=0.5, y=0.5, width=0.4, height=0.4), BoundingBox(x=0.6, y=0.6, width=0.4, height=0.4)
    assert 0 < engine.iou(p, q) < 1
    assert engine.iou(p, q) == pytest.approx(engine.iou(q, p), abs=1e-12)


def test_engine_nms_cases(engine):
    """test_nms_empty_list / single / removes_overlapping / keeps_different_classes / keeps_non_overlapping,
    test_yolo.py:229-313"""
    assert engine.non_max_suppression([], nms_threshold=0.5) == []
    one = [_det(0, 0.9, 0.5, 0.5, 0.2, 0.2)]
    assert engine.non_max_suppression(one, nms_threshold=0.5) == one
    overlapping = [_det(0, 0.9, 0.5, 0.5, 0.3, 0.3), _det(0, 0.8, 0.52, 0.52, 0.3, 0.3)]
    kept = engine.non_max_suppression(overlapping, nms_threshold=0.5)
    assert len(kept) == 1 and kept[0].confidence == 0.9
    two_classes = [_det(0, 0.9, 0.5, 0.5, 0.3, 0.3), _det(1, 0.8, 0.5, 0.5, 0.3, 0.3)]
    assert len(engine.non_max_suppression(two_classes, nms_threshold=0.5)) == 2
    apart = [_det(0, 0.9, 0.2, 0.2, 0.1, 0.1), _det(0, 0.8, 0.8, 0.8, 0.1, 0.1)]
    assert len(engine.non_max_suppression(apart, nms_threshold=0.5)) == 2


def test_deprecated_iou_threshold_keyword_still_works(engine):
    """inference.py:283-296: iou_threshold= warns and is honoured"""
    dets = [_det(0, 0.9, 0.5, 0.5, 0.3, 0.3), _det(0, 0.8, 0.52, 0.52, 0.3, 0.3)]
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        kept = engine.non_max_suppression(dets, iou_threshold=0.5)
    assert len(kept) == 1 and any(issubclass(x.category, DeprecationWarning) for x in w)


def test_detection_string_and_equality():
    """test_detections_list / test_detection_format, test_yolo.py:158-194, 315-334"""
    d = _det(3, 0.75, 0.4, 0.4, 0.2, 0.2)
    assert d == _det(3, 0.75, 0.4, 0.4, 0.2, 0.2) and d != _det(4, 0.75, 0.4, 0.4, 0.2, 0.2)
    assert "class_3" in str(d)
    ranked = sorted([d, _det(1, 0.9, 0.5, 0.5, 0.1, 0.1)], key=lambda x: x.confidence, reverse=True)
    assert ranked[0].class_id == 1


# ------------------------------------------------------------------ pre-NMS validation (inference.py:133-136, 197-208)
class _Fixed(nn.Module):
    """Stands in for the model: returns one fixed (1, S, S, 30) prediction, whatever the image."""

    def __init__(self, pred):
        super().__init__()
        self.S, self.B = 7, 2
        self.register_buffer("pred", pred.clone())

    def forward(self, x):
        return self.pred.unsqueeze(0).expand(x.shape[0], -1, -1, -1)


@pytest.mark.parametrize("case", ["suppressed_x_gt_1", "suppressed_in_range", "two_offenders",
                                  "offender_below_threshold", "confidence_gt_1"])
def test_predict_validates_every_candidate_before_nms(golden, tmp_path, case):
    """tests/golden/validation.npz: what the reference's parse_predictions + non_max_suppression did
    (conf 0.5, nms 0.4).  A candidate outside [0, 1] raises pydantic's ValidationError for the same
    model, field and value even when NMS would have suppressed it -- through predict(), predict_batch()
    and predict_images()."""
    import json

    from PIL import Image
    from pydantic import ValidationError

    g = golden("validation")
    pred = torch.from_numpy(g[f"{case}/pred"])
    want = json.loads(str(g[f"{case}/expect"]))
    eng = YOLOInference(_Fixed(pred), device="cuda")
    path = tmp_path / "grey.png"
    Image.new("RGB", (64, 48), color=(90, 90, 90)).save(path)
    other = torch.zeros(7, 7, 30)
    other[2, 2, 0:5] = torch.tensor([0.5, 0.5, 0.2, 0.2, 0.9])
    other[2, 2, 11] = 1.0
    batch = torch.stack([other, pred, other])
    calls = [lambda: eng.predict(str(path), 0.5, 0.4), lambda: eng.predict_batch(batch, 0.5, 0.4)[1],
             lambda: eng.predict_images([str(path), str(path)], 0.5, 0.4)[1]]
    for call in calls:
        if want["raises"]:
            with pytest.raises(ValidationError) as ei:
                call()
            err = ei.value.errors()[0]
            assert ei.value.title == want["title"]
            assert (err["type"], list(err["loc"]), float(err["input"])) == (want["type"], want["loc"], want["input"])
        else:
            kept = call()
            got = [[d.class_id, d.confidence, d.bbox.x, d.bbox.y, d.bbox.width, d.bbox.height] for d in kept]
            assert got == want["kept"]


def test_lazy_detections_equal_the_eager_lists_and_build_on_demand(golden):
    """predict_batch(lazy=True): SURVEY 8(f) row 2 -- Detection objects are built when read, not when the batch returns;
    the out-of-range check still happens up front."""
    import time

    from pydantic import ValidationError

    from yolo_v1_b200 import synth
    from yolo_v1_b200.inference import LazyDetections

    n = 600
    tgt = synth.make_targets(n, seed=31, p_obj=0.1, size_mix=True)
    pred = synth.make_preds(n, "correlated", seed=31, targets=tgt)
    eng = YOLOInference(_Fixed(pred[0]), device="cuda")
    names = [f"c{i}" for i in range(20)]
    t0 = time.perf_counter()
    eager = eng.predict_batch(pred, 0.05, 0.4, names)
    t1 = time.perf_counter()
    lazy = eng.predict_batch(pred, 0.05, 0.4, names, lazy=True)
    t2 = time.perf_counter()
    assert all(isinstance(d, LazyDetections) for d in lazy) and sum(d.built for d in lazy) == 0
    assert [len(d) for d in lazy] == [len(d) for d in eager] and sum(len(d) for d in eager) > 1000
    i = max(range(n), key=lambda j: len(eager[j]))
    assert lazy[i][0] == eager[i][0] and lazy[i].built == 1 and lazy[i][-1] == eager[i][-1] and lazy[i].built == 2
    assert lazy[i][1:3] == eager[i][1:3] and isinstance(lazy[i][0], Detection) and lazy[i][0].class_name.startswith("c")
    assert (lazy[i].class_ids.tolist(), lazy[i].confidences.tolist()) == ([d.class_id for d in eager[i]], [d.confidence for d in eager[i]])
    with pytest.raises(IndexError):
        lazy[i][len(lazy[i])]
    assert all(a == b for a, b in zip(lazy, eager)) and lazy[i] == eager[i] and not (lazy[i] == eager[i][:-1])
    assert sum(d.built for d in lazy) == sum(len(d) for d in eager)
    assert (t2 - t1) < (t1 - t0)          # nothing was built in the lazy call
    g = golden("validation")
    bad = torch.from_numpy(g["suppressed_x_gt_1/pred"])
    with pytest.raises(ValidationError):
        eng.predict_batch(torch.stack([pred[0], bad]), 0.5, 0.4, lazy=True)


# ------------------------------------------------------------------ evaluate_model (metrics.py:654-714)
def test_evaluate_model_over_a_dataloader_equals_the_metric_fed_the_same_batches():
    """model + DataLoader of HOST tensors + device string in, compute() dict out: the same dict as an mAPMetric
    fed model(images) batch by batch, and as the oracle on the concatenated predictions; model left in eval mode."""
    import numpy as np
    from torch.utils.data import DataLoader, TensorDataset

    from oracle import c_oracle, results
    from yolo_v1_b200 import evaluate_model, synth

    torch.manual_seed(7)
    n = 50                                                  # 4 batches: 16 + 16 + 16 + 2
    tgt = synth.make_targets(n, seed=31, p_obj=0.08, size_mix=True)
    images = torch.randn(n, 3, 32, 32)

    class TinyYolo(nn.Module):
        """A conv stem whose output is mixed with a per-image table so that true positives exist."""

        def __init__(self, table):
            super().__init__()
            self.S, self.B = 7, 2
            self.conv = nn.Conv2d(3, 30, 3, padding=1)
            self.pool = nn.AdaptiveAvgPool2d(7)
            self.drop = nn.Dropout(0.5)                    # would change the output if eval() were not called
            self.register_buffer("table", table)
            self.register_buffer("keys", images.flatten(1).sum(1))

        def forward(self, x):
            idx = (x.flatten(1).sum(1)[:, None] - self.keys[None, :]).abs().argmin(1)
            noise = torch.sigmoid(self.drop(self.pool(self.conv(x)))).permute(0, 2, 3, 1)
            return (0.15 * noise + 0.85 * self.table[idx]).contiguous()

    pred_table = synth.make_preds(n, "correlated", seed=32, targets=tgt)
    model = TinyYolo(pred_table).cuda()
    model.train()
    loader = DataLoader(TensorDataset(images, tgt), batch_size=16, shuffle=False)
    got = evaluate_model(model, loader, "cuda", num_classes=20, conf_threshold=0.05, nms_threshold=0.4)
    assert not model.training                               # metrics.py:697 calls model.eval()

    metric = mAPMetric(num_classes=20, conf_threshold=0.05, nms_threshold=0.4)
    outs = []
    with torch.no_grad():
        for x, t in loader:
            out = model(x.cuda())
            outs.append(out.cpu())
            metric.update(out, t.cuda())
    want = metric.compute()
    assert got == want and float(got["mAP50"]) > 0.0
    allp = torch.cat(outs)
    st = c_oracle.update_batch(allp.numpy(), tgt.numpy(), 0.05, 0.4)
    ref = c_oracle.compute(st, metric.iou_thresholds)
    exp = results.assemble(ref["ap"], ref["overall"], ref["sizes"], metric.iou_thresholds, 20, n)
    assert sorted(got) == sorted(exp)
    for k in exp:
        assert float(got[k]) == pytest.approx(float(exp[k]), rel=1e-9, abs=1e-12), k
