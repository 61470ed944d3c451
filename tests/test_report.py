"""yolo_v1_b200.report against the text the reference's own report code produced
(tests/golden/report.json, made by tests/golden/make_report_golden.py from the unmodified
src/evaluate.py:185-290 and src/predict.py:94-108).  Host-side formatting: runs without a GPU."""
import json
import os

import pytest

from yolo_v1_b200 import report
from yolo_v1_b200.schemas import BoundingBox, Detection

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "report.json")))


@pytest.mark.parametrize("case", ["corr64", "ties48"])
def test_evaluation_report_matches_reference_text(case, tmp_path):
    g = GOLD[case]
    res = dict(g["results"])
    ckpt = str(tmp_path / "best.pth")
    printed = report.format_evaluation_report(res, num_classes=20)
    path = report.write_results_file(res, ckpt, "VOC2007 test", g["n_images"])
    assert path == tmp_path / "evaluation_results.txt"
    want_printed = g["printed"].replace("<DIR>", str(tmp_path))
    assert printed + f"\nResults saved to {path}\n" == want_printed
    assert path.read_text() == g["written"].replace("<DIR>", str(tmp_path))
    assert report.format_results_file(res, ckpt, "VOC2007 test", g["n_images"]) == path.read_text()


def test_class_table_is_sorted_by_ap_with_stable_ties():
    res = {f"AP50:95_class_{i}": v for i, v in enumerate([0.2, 0.5, 0.2, 0.9])}
    assert [c for c, _ in report.classes_by_ap(res, 4)] == [3, 1, 0, 2]


def test_missing_ap75_key_raises_like_the_reference():
    res = dict(GOLD["corr64"]["results"])
    del res["mAP75"]   # a threshold list without 0.75: evaluate.py fails with KeyError on this line too
    with pytest.raises(KeyError):
        report.format_evaluation_report(res)


def test_detection_lines_match_predict_py():
    g = GOLD["predict"]
    dets = [Detection(class_id=c, class_name=n, confidence=cf, bbox=BoundingBox(x=x, y=y, width=w, height=h))
            for c, n, cf, x, y, w, h in g["detections"]]
    assert report.format_detections("img/dog.jpg", dets) == g["printed"]


def test_voc_class_names_are_the_references():
    assert len(report.VOC_CLASSES) == 20 and report.VOC_CLASSES[0] == "aeroplane" and report.VOC_CLASSES[-1] == "tvmonitor"
    table = GOLD["corr64"]["printed"]
    assert all(name in table for name in report.VOC_CLASSES)
