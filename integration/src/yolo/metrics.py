"""Replaces src/yolo/metrics.py of mattiaskvist/yolo-v1: `from ..metrics import evaluate_model`
(src/yolo/training/trainer.py:12), `from src.yolo.metrics import mAPMetric` (tests/test_metrics.py:7) and
`from .metrics import mAPMetric, evaluate_model` (src/yolo/__init__.py:5) now get the CUDA metric."""
from yolo_v1_b200.metrics import evaluate_model, mAPMetric  # noqa: F401

__all__ = ["evaluate_model", "mAPMetric"]
