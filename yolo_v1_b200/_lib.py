"""ctypes binding of libyolohead.so (include/yolohead.h).

There is no fallback: if the library is missing or a call fails, the caller gets
an exception.  Nothing here touches oracle/.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("YH_LIB") or os.path.join(_HERE, "libyolohead.so")   # YH_LIB: an experiment build

F32, F16, BF16 = 0, 1, 2
IOU_METRICS, IOU_INFER = 0, 1
ORDER_CLASS_GROUPED, ORDER_GLOBAL_CONF = 0, 1
MAX_THRESHOLDS = 13

_DTYPES = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}

_vp, _i, _d, _f, _sz, _i64 = C.c_void_p, C.c_int, C.c_double, C.c_float, C.c_size_t, C.c_int64

# name -> (restype, argtypes); must list every symbol include/yolohead.h declares
SIGNATURES = {
    "yh_version": (_i, []),
    "yh_last_error": (C.c_char_p, []),
    "yh_launch_count": (C.c_longlong, []),
    "yh_loss_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "yh_loss_fwd_bwd": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _f, _f, _vp, _vp, _vp, _sz, _vp]),
    "yh_scale_inplace": (_i, [_vp, _i, _sz, _vp, _vp]),
    "yh_loss_static_scratch_bytes": (_sz, []),
    "yh_loss_fwd_bwd_static": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _f, _f, _vp, _vp, _vp]),
    "yh_head_linear_loss_workspace_bytes": (_sz, []),
    "yh_head_linear_loss": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _f, _f, _f, _i, _vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "yh_decode_nms": (_i, [_vp, _i, _i, _i, _i, _d, _d, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "yh_map_update": (_i, [_vp, _vp, _i, _i, _i, _i, _d, _d, _vp, _i] + [_vp] * 11),
    "yh_map_update_loss_workspace_bytes": (_sz, [_i]),
    "yh_map_update_loss": (_i, [_vp, _vp, _i, _i, _i, _i, _d, _d, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp,
                                _vp, _sz, _i, _vp]),
    "yh_list_form": (_i, [_vp] * 8 + [_i, _i, _i, _i] + [_vp] * 4),
    "yh_parse_gt": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "yh_compact_workspace_bytes": (_sz, [_i]),
    "yh_compact_records": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "yh_ap_workspace_bytes": (_sz, [_i64, _i, _i]),
    "yh_ap_reduce": (_i, [_vp, _vp, _i64, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "yh_preprocess_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "yh_preprocess_rgb8": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "yh_partition_workspace_bytes": (_sz, [_i64]),
    "yh_partition_records": (_i, [_vp, _vp, _i64, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "yh_peer_sum": (_i, [_vp, _i, _i, _i, _i, _vp, C.c_uint, _vp]),
    "yh_peer_alloc": (_i, [_sz, _vp, _vp]),
    "yh_peer_open": (_i, [_vp, _vp]),
    "yh_peer_set_wait_ms": (_i, [_vp, C.c_uint]),
    "yh_peer_close": (_i, [_vp]),
    "yh_peer_free": (_i, [_vp]),
    "yh_peer_block_bytes": (_sz, [_i64]),
    "yh_peer_key_offset": (_sz, [_i64]),
    "yh_peer_flags_offset": (_sz, [_i64]),
    "yh_exchange_workspace_bytes": (_sz, [_i64]),
    "yh_exchange_records": (_i, [_vp, _vp, _i64, _vp, _vp, _i, _i, _i, _vp, _i64, C.c_uint, _vp, _i64, _vp, _vp,
                                 _vp, _sz, _vp]),
    "yh_exchange_images_workspace_bytes": (_sz, [_i64, _i]),
    "yh_exchange_images": (_i, [_i, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _vp, _i64, C.c_uint, _vp, _vp, _vp, _vp, _sz,
                                _vp]),
    "yh_encode_targets": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "yh_iou_pairs": (_i, [_vp, _vp, _i64, _i, _vp, _vp]),
    "yh_nms_lists": (_i, [_vp, _i, _vp, _vp, _vp, _d, _i, _i, _vp, _vp, _i, _vp]),
    "yh_match_lists": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _i, _vp, _vp, _vp, _i, _vp]),
}

_lib = None


class YoloHeadError(RuntimeError):
    pass


def lib():
    """The loaded library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise YoloHeadError(
                f"{LIB_PATH} not found: build it with `python -m yolo_v1_b200.build` "
                "(the head hot path has no CPU or PyTorch fallback)")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().yh_last_error().decode("utf-8", "replace")
        kind = {-1: "bad argument", -2: "unsupported shape", -3: "workspace too small", -4: "misaligned pointer"}.get(
            rc, f"CUDA error {rc}" if rc > 0 else f"error {rc}")
        if rc == -2:
            raise NotImplementedError(f"{what}: {msg}")
        raise YoloHeadError(f"{what}: {kind}: {msg}")


def dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported prediction dtype {t.dtype}; use float32, float16 or bfloat16") from None


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise YoloHeadError(
            f"{name} is on {t.device}: the YOLOv1 head kernels run on CUDA only (no CPU fallback). "
            "Move the tensor to the GPU first.")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device) -> int:
    """cudaStream_t of torch's current stream on `device` (the raw-handle call skips building a Stream object)."""
    if _raw_stream is not None:
        idx = device.index if isinstance(device, torch.device) else torch.device(device).index
        return _raw_stream(torch.cuda.current_device() if idx is None else idx)
    return torch.cuda.current_stream(device).cuda_stream


class _NoSwitch:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NO_SWITCH = _NoSwitch()


def on_device(device):
    """Context in which `device` is the current CUDA device: free when it already is."""
    idx = device.index if isinstance(device, torch.device) else torch.device(device).index
    if idx is None or idx == torch.cuda.current_device():
        return _NO_SWITCH
    return torch.cuda.device(idx)


def ptr(t):
    return None if t is None else t.data_ptr()


class Workspace:
    """Scratch buffers for kernels that need one, one per (device, stream), grown on demand.

    Keyed by stream because a workspace carries per-launch state (sort control words, the loss
    kernel's partials, exchange staging): two calls in flight on different streams -- or an eager
    call next to a replaying CUDA graph, which owns a private stream -- must not share one.
    While a `holding()` context is open every buffer handed out is also appended to the holder's
    list: CapturedHeadStep keeps the buffers whose addresses its graph has baked in alive that way,
    whatever later calls do to this cache."""

    _holders: list = []

    def __init__(self, zeroed: bool = False, align: int = 256):
        self._buf: dict = {}
        self._zeroed, self._align = zeroed, align

    def get(self, device, nbytes: int, stream: int | None = None) -> torch.Tensor:
        device = torch.device(device)
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        key = (device.index, stream_ptr(device) if stream is None else stream)
        buf = self._buf.get(key)
        if buf is None or buf.numel() < nbytes + self._align:
            make = torch.zeros if self._zeroed else torch.empty
            buf = make(max(nbytes, 1 << 16) + self._align, dtype=torch.uint8, device=device)
            self._buf[key] = buf
        for h in Workspace._holders:
            h.append(buf)
        pad = (-buf.data_ptr()) % self._align
        return buf[pad:pad + nbytes]

    @staticmethod
    def holding(refs: list):
        return _Holding(refs)


class _Holding:
    def __init__(self, refs):
        self.refs = refs

    def __enter__(self):
        Workspace._holders.append(self.refs)
        return self.refs

    def __exit__(self, *a):
        Workspace._holders.remove(self.refs)
        return False
