"""Replaces src/yolo/loss.py of mattiaskvist/yolo-v1: `from yolo.loss import YOLOLoss` (src/train.py:15) and
`from ..loss import YOLOLoss` (src/yolo/training/trainer.py:11) now get the fused CUDA loss."""
from yolo_v1_b200.loss import YOLOLoss  # noqa: F401

__all__ = ["YOLOLoss"]
