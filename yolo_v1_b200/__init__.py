"""yolo_v1_b200 -- B200-native (sm_100a) implementation of the YOLOv1 detection-head hot path.

Drop-in counterparts of the reference's loss / metrics / inference modules
(mattiaskvist/yolo-v1, src/yolo/{loss,metrics,inference,schemas}.py) over
hand-written CUDA kernels behind a C ABI (include/yolohead.h, libyolohead.so).
The conv backbone stays stock PyTorch/cuDNN and is not part of this package.
"""
from .graph import CapturedHeadStep
from .head_linear import FusedHeadLoss, head_linear
from .inference import YOLOInference
from .loss import YOLOLoss
from .metrics import evaluate_model, mAPMetric
from .report import format_evaluation_report, write_results_file
from .schemas import BoundingBox, Detection
from .targets import encode_target, encode_targets, encode_targets_csr

__version__ = "0.1.0"

__all__ = ["BoundingBox", "CapturedHeadStep", "Detection", "FusedHeadLoss", "YOLOInference", "YOLOLoss", "encode_target", "encode_targets",
           "encode_targets_csr", "evaluate_model", "format_evaluation_report", "head_linear", "mAPMetric", "write_results_file"]
