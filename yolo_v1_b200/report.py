"""Evaluation report of the reference's evaluate.py (src/evaluate.py:185-290) for a results dict
as mAPMetric.compute() / evaluate_model() return it.

The reference prints the report and writes `evaluation_results.txt` from inside its CLI `main()`;
here the two texts are functions of the dict so a caller of `evaluate_model` (or of
`distributed.evaluate_sharded`, on rank 0) gets the same output without the CLI.  Host-side string
formatting only: nothing here touches the GPU.  SURVEY.md section 8f row 1.
"""
from __future__ import annotations

from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

# VOCDetectionYOLO.VOC_CLASSES (src/yolo/dataset.py:52-73), the names evaluate.py prints
VOC_CLASSES = [
    "aeroplane", "bicycle", "bird", "boat", "bottle", "bus", "car", "cat", "chair", "cow", "diningtable", "dog",
    "horse", "motorbike", "person", "pottedplant", "sheep", "sofa", "train", "tvmonitor",
]

_RULE = "=" * 70
_DASH = "-" * 70
_SIZES = (("small", "  Small objects  (area < 32x32):"),
          ("medium", "  Medium objects (32x32 <= area < 96x96):"),
          ("large", "  Large objects  (area >= 96x96):"))


def classes_by_ap(results: Dict[str, float], num_classes: int) -> List[Tuple[int, float]]:
    """(class id, AP50:95) sorted by AP50:95 descending; list.sort is stable, so ties keep class order
    (evaluate.py:225-228)."""
    aps = [(i, results[f"AP50:95_class_{i}"]) for i in range(num_classes)]
    aps.sort(key=lambda x: x[1], reverse=True)
    return aps


def _summary(results: Dict[str, float]) -> List[str]:
    return [f"mAP50:95: {results['mAP50:95']:.4f}", f"mAP@0.5: {results['mAP50']:.4f}",
            f"mAP@0.75: {results['mAP75']:.4f}", f"Precision: {results['precision']:.4f}",
            f"Recall: {results['recall']:.4f}"]


def _size_block(results: Dict[str, float]) -> List[str]:
    out = []
    for key, title in _SIZES:
        out += [title,
                f"    mAP50:95: {results[f'mAP50:95_{key}']:.4f}  (n={results[f'num_{key}_objects']})",
                f"    mAP@0.5:  {results[f'mAP50_{key}']:.4f}",
                f"    mAP@0.75: {results[f'mAP75_{key}']:.4f}"]
    return out


def _class_table(results: Dict[str, float], num_classes: int, class_names: Sequence[str]) -> List[str]:
    out = [f"{'Class':<15} {'AP50:95':>8} {'AP50':>8} {'AP75':>8}", _DASH]
    for class_id, ap_5095 in classes_by_ap(results, num_classes):
        out.append(f"{class_names[class_id]:<15} {ap_5095:>8.4f} {results[f'AP50_class_{class_id}']:>8.4f} "
                   f"{results[f'AP75_class_{class_id}']:>8.4f}")
    return out


def format_evaluation_report(results: Dict[str, float], num_classes: int = 20,
                             class_names: Optional[Sequence[str]] = None) -> str:
    """The text evaluate.py prints between "EVALUATION RESULTS" and the closing rule
    (evaluate.py:185-241), newline-terminated, exactly as its print() calls emit it.
    Needs the default thresholds (the mAP75 / AP75 keys exist only when 0.75 is one of them), like
    the reference, which raises KeyError otherwise."""
    names = VOC_CLASSES if class_names is None else class_names
    lines = ["", _RULE, "EVALUATION RESULTS", _RULE] + _summary(results) + [""]
    lines += ["Size-Based Metrics:", _DASH] + _size_block(results) + [""]
    lines += ["Per-class Average Precision:", _DASH] + _class_table(results, num_classes, names) + [_RULE]
    return "\n".join(lines) + "\n"


def format_results_file(results: Dict[str, float], checkpoint: str, dataset_info: str, num_images: int,
                        num_classes: int = 20, class_names: Optional[Sequence[str]] = None) -> str:
    """Contents of evaluation_results.txt (evaluate.py:252-288)."""
    names = VOC_CLASSES if class_names is None else class_names
    lines = ["YOLO v1 Evaluation Results", _RULE, f"Checkpoint: {checkpoint}", f"Dataset: {dataset_info}",
             f"Number of images: {num_images}", ""] + _summary(results)
    lines += ["", "Size-Based Metrics:", _DASH] + _size_block(results)
    lines += ["", "Per-class Average Precision:", _DASH] + _class_table(results, num_classes, names) + [_RULE]
    return "\n".join(lines) + "\n"


def write_results_file(results: Dict[str, float], checkpoint: str, dataset_info: str, num_images: int,
                       num_classes: int = 20, class_names: Optional[Sequence[str]] = None) -> Path:
    """Writes evaluation_results.txt next to the checkpoint, where evaluate.py puts it (evaluate.py:244)."""
    path = Path(checkpoint).parent / "evaluation_results.txt"
    path.write_text(format_results_file(results, checkpoint, dataset_info, num_images, num_classes, class_names))
    return path


def format_detections(image_path, detections) -> str:
    """What predict.py prints per image (predict.py:94-108)."""
    lines = [f"\nProcessing: {image_path}", f"Found {len(detections)} objects:"]
    lines += [f"  {i}. {det.class_name}: {det.confidence:.2f} at ({det.bbox.x:.2f}, {det.bbox.y:.2f}) "
              f"size ({det.bbox.width:.2f}, {det.bbox.height:.2f})" for i, det in enumerate(detections, 1)]
    return "\n".join(lines) + "\n"
