"""Image preprocessing of YOLOInference on the GPU (yh_preprocess_rgb8).

`preprocess_rgb8` is `transforms.Compose([Resize((448, 448)), ToTensor(), Normalize(mean, std)])(image)` of the
reference (src/yolo/inference.py:57-66) for an RGB uint8 image, bit for bit: Pillow's fixed-point bilinear
resample and torchvision's float32 scaling, computed by three small kernels instead of on the host.
SURVEY.md section 8f row 2.
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import numpy as np
import torch

from . import _lib

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)
_WS = _lib.Workspace()


def _workspace(dev: torch.device, nbytes: int) -> torch.Tensor:
    return _WS.get(dev, nbytes)


def preprocess_rgb8(image: torch.Tensor, size: Sequence[int] = (448, 448), mean: Sequence[float] = IMAGENET_MEAN,
                    std: Sequence[float] = IMAGENET_STD) -> torch.Tensor:
    """image: (H, W, 3) uint8 CUDA tensor (the bytes of a PIL RGB image) -> (3, size[0], size[1]) float32;
    or a batch (N, H, W, 3) of equal-sized images -> (N, 3, size[0], size[1]), three launches for all of them."""
    _lib.require_cuda(image, "image")
    if image.dtype is not torch.uint8 or image.dim() not in (3, 4) or image.shape[-1] != 3:
        raise TypeError("preprocess_rgb8 wants an (H, W, 3) or (N, H, W, 3) uint8 tensor")
    single = image.dim() == 3
    image = image.contiguous()
    if image.data_ptr() % 4:
        image = image.clone()
    dev = image.device
    N = 1 if single else int(image.shape[0])
    H, W = int(image.shape[-3]), int(image.shape[-2])
    oh, ow = int(size[0]), int(size[1])
    if N == 0:
        return torch.empty(0, 3, oh, ow, dtype=torch.float32, device=dev)
    lib = _lib.lib()
    nbytes = lib.yh_preprocess_workspace_bytes(N, H, W, oh, ow)
    if nbytes == 0:
        raise ValueError(f"preprocess_rgb8: unsupported shape {N}x{H}x{W} -> {oh}x{ow}")
    ws = _workspace(dev, nbytes)
    out = torch.empty(N, 3, oh, ow, dtype=torch.float32, device=dev)
    m = (C.c_float * 3)(*[float(np.float32(v)) for v in mean])
    s = (C.c_float * 3)(*[float(np.float32(v)) for v in std])
    with _lib.on_device(dev):
        rc = lib.yh_preprocess_rgb8(image.data_ptr(), N, H, W, oh, ow, C.cast(m, C.c_void_p), C.cast(s, C.c_void_p),
                                    out.data_ptr(), ws.data_ptr(), nbytes, _lib.stream_ptr(dev))
    _lib.check(rc, "yh_preprocess_rgb8")
    return out[0] if single else out


def preprocess_pil(image, device, size: Sequence[int] = (448, 448)) -> torch.Tensor:
    """A PIL image (any mode; converted to RGB like load_image does) -> (3, H', W') float32 on `device` (CUDA)."""
    arr = np.array(image.convert("RGB") if image.mode != "RGB" else image)  # a writable copy of the pixel bytes
    dev = torch.device(device)
    return preprocess_rgb8(torch.from_numpy(np.ascontiguousarray(arr)).to(dev, non_blocking=True), size)
