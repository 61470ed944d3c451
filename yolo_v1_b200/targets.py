"""Target encoding on the GPU -- batched counterpart of VOCDetectionYOLO._encode_target
(/root/reference/src/yolo/dataset.py:487-532).

The reference builds one (S, S, B*5 + C) target per sample inside the DataLoader, in
Python.  `encode_targets` does a whole batch in one launch of yh_encode_targets from the
normalised boxes and class ids the dataset already extracts
(`_extract_bboxes_from_annotation`, dataset.py:430-471): same cell assignment, same
"first box of a cell wins" rule, same cell-relative coordinates rounded to fp32.
"""
from __future__ import annotations

from typing import Sequence

import torch

from . import _lib


def encode_targets(bboxes: Sequence[Sequence[Sequence[float]]], class_ids: Sequence[Sequence[int]],
                   S: int = 7, B: int = 2, C: int = 20, device=None) -> torch.Tensor:
    """bboxes[n] = [[x_center, y_center, w, h], ...] (normalised), class_ids[n] = [int, ...]
    -> float32 tensor (N, S, S, B*5 + C) on `device`.

    Raises IndexError / ValueError where the reference's loop would (a centre that maps
    outside the grid after Python's negative-index wrap, a non-finite centre, a class id
    outside the one-hot range)."""
    if len(bboxes) != len(class_ids):
        raise ValueError("bboxes and class_ids differ in length")
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    if dev.type != "cuda":
        raise _lib.YoloHeadError("encode_targets runs on CUDA only (no CPU fallback)")
    n = len(bboxes)
    counts = [min(len(b), len(c)) for b, c in zip(bboxes, class_ids)]  # the reference zips the two lists
    off = [0]
    for m in counts:
        off.append(off[-1] + m)
    flat_b = [list(map(float, box)) for b, m in zip(bboxes, counts) for box in b[:m]]
    flat_c = [int(c) for cs, m in zip(class_ids, counts) for c in cs[:m]]
    boxes_d = torch.tensor(flat_b if flat_b else [[0.0] * 4], dtype=torch.float64).reshape(-1, 4).to(dev)
    cls_d = torch.tensor(flat_c if flat_c else [0], dtype=torch.int32).to(dev)
    off_d = torch.tensor(off, dtype=torch.int64).to(dev)
    out, bad = encode_targets_csr(boxes_d, cls_d, off_d, S, B, C)
    if n and bool(bad.any()):
        which = int(torch.nonzero(bad)[0])
        raise IndexError(f"sample {which}: a box centre or class id falls outside the target tensor "
                         "(the reference's _encode_target raises here too)")
    return out


def encode_targets_csr(boxes: torch.Tensor, class_ids: torch.Tensor, offsets: torch.Tensor, S: int = 7, B: int = 2,
                       C: int = 20):
    """Tensor form for loaders that keep a batch's boxes flat on the device: boxes (M, 4) float64
    (x_center, y_center, w, h), class_ids (M,) int32, offsets (N + 1,) int64 with image n owning
    rows [offsets[n], offsets[n + 1]).  Returns (targets (N, S, S, B*5 + C) float32, bad (N,) uint8)
    without reading anything back; bad[n] = 1 where the reference would raise for sample n."""
    for t, name in ((boxes, "boxes"), (class_ids, "class_ids"), (offsets, "offsets")):
        _lib.require_cuda(t, name)
    if boxes.dtype is not torch.float64 or class_ids.dtype is not torch.int32 or offsets.dtype is not torch.int64:
        raise TypeError("encode_targets_csr wants float64 boxes, int32 class_ids and int64 offsets")
    dev = boxes.device
    n = offsets.numel() - 1
    if n < 0:
        raise ValueError("offsets must hold N + 1 entries")
    boxes, class_ids, offsets = boxes.contiguous(), class_ids.contiguous(), offsets.contiguous()
    out = torch.empty(n, S, S, B * 5 + C, dtype=torch.float32, device=dev)
    bad = torch.zeros(max(n, 1), dtype=torch.uint8, device=dev)
    with _lib.on_device(dev):
        rc = _lib.lib().yh_encode_targets(boxes.data_ptr(), class_ids.data_ptr(), offsets.data_ptr(), n, S, B, C,
                                          out.data_ptr(), bad.data_ptr(), _lib.stream_ptr(dev))
    _lib.check(rc, "yh_encode_targets")
    return out, bad[:n]


def encode_target(bboxes: Sequence[Sequence[float]], class_ids: Sequence[int], S: int = 7, B: int = 2, C: int = 20,
                  device=None) -> torch.Tensor:
    """Single-sample form with _encode_target's own signature: -> (S, S, B*5 + C)."""
    return encode_targets([bboxes], [class_ids], S, B, C, device)[0]
