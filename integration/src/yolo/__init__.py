"""Stand-in for src/yolo/__init__.py of mattiaskvist/yolo-v1, reduced to the hot-path names.

In the reference tree this file stays AS IT IS (src/yolo/__init__.py:3-13 imports `.loss`, `.metrics`,
`.schemas` by module path); only the four modules next to it -- loss.py, metrics.py, inference.py,
schemas.py -- are replaced by the re-export shims in this directory.  This stand-in exists so that the
shims can be imported (`yolo.loss`, `src.yolo.metrics`, ...) where the rest of the reference package
(models, dataset, training, utils) is not present, e.g. on the GPU test box."""
from .loss import YOLOLoss
from .metrics import evaluate_model, mAPMetric
from .schemas import BoundingBox, Detection

__all__ = ["BoundingBox", "Detection", "YOLOLoss", "evaluate_model", "mAPMetric"]
