"""The binding a maintainer adds: integration/src/yolo/{loss,metrics,inference,
schemas}.py replace the reference's files of the same name.  Every import spelling the reference uses must
resolve to the CUDA drop-in."""
import importlib
import os
import shutil
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIMS = os.path.join(ROOT, "integration", "src", "yolo")
REF = "/root/reference/src/yolo"


def _is_ours(k):
    return k in ("yolo", "src") or k.startswith(("yolo.", "src."))


def _fresh_import(paths, names):
    """Import `names` with `paths` in front of sys.path and no cached `yolo` / `src` modules."""
    saved_path, saved_mods = list(sys.path), {k: v for k, v in sys.modules.items() if _is_ours(k)}
    for k in saved_mods:
        del sys.modules[k]
    sys.path[:0] = paths
    try:
        return [importlib.import_module(n) for n in names]
    finally:
        sys.path[:] = saved_path
        for k in [k for k in sys.modules if _is_ours(k)]:
            del sys.modules[k]
        sys.modules.update(saved_mods)


def test_every_reference_import_spelling_resolves_to_the_drop_in():
    import yolo_v1_b200 as y

    loss, metrics, inference, schemas, pkg = _fresh_import(
        [os.path.join(ROOT, "integration", "src")], ["yolo.loss", "yolo.metrics", "yolo.inference", "yolo.schemas", "yolo"])
    assert loss.YOLOLoss is y.YOLOLoss                                    # src/train.py:15
    assert metrics.mAPMetric is y.mAPMetric and metrics.evaluate_model is y.evaluate_model
    assert inference.YOLOInference is y.YOLOInference                     # src/predict.py:19
    assert schemas.Detection is y.Detection and schemas.BoundingBox is y.BoundingBox
    assert pkg.YOLOLoss is y.YOLOLoss and pkg.evaluate_model is y.evaluate_model
    (m2,) = _fresh_import([os.path.join(ROOT, "integration")], ["src.yolo.metrics"])   # tests/test_metrics.py:7
    assert m2.mAPMetric is y.mAPMetric


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference tree (build container only)")
def test_shims_overlaid_on_the_reference_package_rewire_its_own_modules(tmp_path):
    """A temporary copy of the reference's src/yolo with the four files replaced: the reference's OWN
    __init__.py and training/trainer.py (relative imports `..loss`, `..metrics`) now bind the drop-in."""
    import yolo_v1_b200 as y

    dst = tmp_path / "src" / "yolo"
    shutil.copytree(REF, dst, ignore=shutil.ignore_patterns("__pycache__"))
    for f in ("loss.py", "metrics.py", "inference.py", "schemas.py"):
        shutil.copy(os.path.join(SHIMS, f), dst / f)
    pkg, trainer = _fresh_import([str(tmp_path / "src")], ["yolo", "yolo.training.trainer"])
    assert pkg.__file__.startswith(str(tmp_path))
    assert pkg.YOLOLoss is y.YOLOLoss and pkg.mAPMetric is y.mAPMetric and pkg.Detection is y.Detection
    assert trainer.YOLOLoss is y.YOLOLoss and trainer.evaluate_model is y.evaluate_model   # trainer.py:11-12
    assert hasattr(pkg, "YOLOv1") and hasattr(pkg, "ResNetBackbone")      # the rest of the package is the reference's


@pytest.mark.gpu
def test_one_step_through_the_shim_spelling(tmp_path):
    """criterion(...), metric.update/compute and engine.predict, imported the way the reference's scripts do."""
    from PIL import Image

    from yolo_v1_b200 import synth

    loss, metrics, inference = _fresh_import([os.path.join(ROOT, "integration", "src")],
                                             ["yolo.loss", "yolo.metrics", "yolo.inference"])
    tgt = synth.make_targets(16, seed=5, p_obj=0.1, size_mix=True)
    pred = synth.make_preds(16, "correlated", seed=6, targets=tgt)
    p = pred.cuda().requires_grad_(True)
    total, parts = loss.YOLOLoss()(p, tgt.cuda())
    total.backward()
    assert set(parts) == {"total", "coord", "conf_obj", "conf_noobj", "class"} and p.grad.abs().sum() > 0
    m = metrics.mAPMetric(num_classes=20)
    m.update(pred, tgt)                       # host tensors, as tests/test_metrics.py hands them over
    res = m.compute()
    assert 0.0 < float(res["mAP50"]) <= 1.0

    class Head(torch.nn.Module):
        S, B = 7, 2

        def forward(self, x):
            return pred[:1].to(x.device).expand(x.shape[0], -1, -1, -1).clamp(0, 1)

    path = tmp_path / "img.png"
    Image.new("RGB", (80, 60), color=(10, 200, 30)).save(path)
    dets = inference.YOLOInference(Head(), device="cuda").predict(str(path), conf_threshold=0.2)
    assert isinstance(dets, list)
