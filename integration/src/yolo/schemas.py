"""Replaces src/yolo/schemas.py of mattiaskvist/yolo-v1 (`from yolo.schemas import Detection, BoundingBox`,
tests/test_yolo.py:10; `from .schemas import Detection, BoundingBox`, src/yolo/__init__.py:13): the same
pydantic models, shared with the CUDA inference engine so that `isinstance` checks agree."""
from yolo_v1_b200.schemas import BoundingBox, Detection  # noqa: F401

__all__ = ["BoundingBox", "Detection"]
