"""Replaces src/yolo/inference.py of mattiaskvist/yolo-v1: `from yolo.inference import YOLOInference`
(src/predict.py:19, tests/test_yolo.py:8) now gets the engine whose decode + NMS run on CUDA."""
from yolo_v1_b200.inference import YOLOInference  # noqa: F401

__all__ = ["YOLOInference"]
