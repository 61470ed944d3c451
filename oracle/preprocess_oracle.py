"""CPU restatement of the image preprocessing of YOLOInference (test infrastructure: only tests/,
__graft_entry__.smoke() and bench.py's CPU legs may import this).

Reference: src/yolo/inference.py:57-66,88-101 -- torchvision `Resize((448, 448))` on a PIL image, `ToTensor()`,
`Normalize(mean, std)`.  The arithmetic lives in third-party code that is not under /root/reference:
  * Pillow (uv.lock pins 11.3.0; 12.2.0 is installed here), `Image.resize(size, BILINEAR)` ->
    src/libImaging/Resample.c: `precompute_coeffs` (double), `normalize_coeffs_8bpc` (22-bit fixed point),
    `ImagingResampleHorizontal_8bpc` then `ImagingResampleVertical_8bpc` with an 8-bit intermediate image;
  * torchvision 0.23 (0.26 here): `to_tensor` = uint8 -> float32, `.div(255)`; `normalize` = `.sub_(mean).div_(std)`
    with float32 mean / std.
Pinned against the real libraries in this container by tests/test_preprocess.py (random images of many
sizes, bit for bit) and against tests/golden/preprocess.npz.
"""
import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2
MEAN = (0.485, 0.456, 0.406)
STD = (0.229, 0.224, 0.225)


def _bilinear(x: float) -> float:
    if x < 0.0:
        x = -x
    return 1.0 - x if x < 1.0 else 0.0


def precompute_coeffs(in_size: int, out_size: int):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the box (0, in_size) and the bilinear filter
    (support 1.0).  Returns (bounds[out_size][2] = (first input index, count), kk[out_size][ksize] int32)."""
    scale = filterscale = float(in_size) / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        ww = 0.0
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        k = [0.0] * ksize
        for x in range(xmax):
            w = _bilinear((x + xmin - center + 0.5) * ss)
            k[x] = w
            ww += w
        for x in range(xmax):
            if ww != 0.0:
                k[x] /= ww
        bounds[xx] = (xmin, xmax)
        for x in range(ksize):
            v = k[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
    return bounds, kk


def _resample_axis0(img: np.ndarray, out_size: int) -> np.ndarray:
    """One pass along axis 0 of a uint8 array (n, ...) -> (out_size, ...)."""
    bounds, kk = precompute_coeffs(img.shape[0], out_size)
    out = np.empty((out_size,) + img.shape[1:], np.uint8)
    src = img.astype(np.int64)
    for xx in range(out_size):
        x0, n = int(bounds[xx, 0]), int(bounds[xx, 1])
        acc = np.full(img.shape[1:], 1 << (PRECISION_BITS - 1), np.int64)
        for x in range(n):
            acc += src[x0 + x] * int(kk[xx, x])
        out[xx] = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return out


def resize_bilinear_u8(img: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """PIL `Image.resize((out_w, out_h), BILINEAR)` of an (H, W, 3) uint8 image: horizontal pass first (skipped when
    the width already matches), then vertical (skipped when the height matches), 8-bit rounding in between."""
    h, w = img.shape[:2]
    if w != out_w:
        img = np.ascontiguousarray(_resample_axis0(np.ascontiguousarray(img.transpose(1, 0, 2)), out_w).transpose(1, 0, 2))
    if h != out_h:
        img = _resample_axis0(img, out_h)
    return img


def normalize_chw(img_u8: np.ndarray, mean=MEAN, std=STD) -> np.ndarray:
    """ToTensor + Normalize: (H, W, 3) uint8 -> (3, H, W) float32, every step rounded to float32."""
    x = img_u8.transpose(2, 0, 1).astype(np.float32) / np.float32(255)
    m = np.asarray(mean, np.float32).reshape(3, 1, 1)
    s = np.asarray(std, np.float32).reshape(3, 1, 1)
    return ((x - m) / s).astype(np.float32)


def preprocess(img_u8: np.ndarray, out_h: int = 448, out_w: int = 448) -> np.ndarray:
    return normalize_chw(resize_bilinear_u8(img_u8, out_h, out_w))
