"""numpy restatement of VOCDetectionYOLO._encode_target (CPU checker for yh_encode_targets).

TEST INFRASTRUCTURE ONLY -- never imported by yolo_v1_b200/.

Follows /root/reference/src/yolo/dataset.py:
  zero target (S, S, 5B + C)                       dataset.py:504
  i = int(S*y), j = int(S*x), clamped to S-1        dataset.py:509-515
  x_cell = S*x - j, y_cell = S*y - i                dataset.py:517-518
  first box of a cell wins (target[i,j,4] == 0)     dataset.py:521
  box, confidence 1, class one-hot                  dataset.py:523-530
Python doubles throughout, rounded to float32 on assignment like the tensor store.
Pinned against the reference's own outputs in tests/golden/encode.npz
(tests/test_oracle_golden.py).
"""
from __future__ import annotations

import numpy as np


def encode_target(bboxes, class_ids, S=7, B=2, C=20) -> np.ndarray:
    target = np.zeros((S, S, 5 * B + C), dtype=np.float32)
    for bbox, class_id in zip(bboxes, class_ids):
        x_center, y_center, width, height = (float(v) for v in bbox)
        i = int(S * y_center)
        j = int(S * x_center)
        i = min(i, S - 1)
        j = min(j, S - 1)
        x_cell = S * x_center - j
        y_cell = S * y_center - i
        if target[i, j, 4] == 0:
            target[i, j, 0] = x_cell
            target[i, j, 1] = y_cell
            target[i, j, 2] = width
            target[i, j, 3] = height
            target[i, j, 4] = 1.0
            target[i, j, 5 * B + int(class_id)] = 1.0
    return target
