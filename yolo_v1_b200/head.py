"""Functional wrappers over the C ABI (include/yolohead.h): tensors in, tensors out.

Everything here enqueues CUDA kernels on the current stream of the input's device
and returns device tensors; nothing synchronises unless the caller reads a result.
No CPU or PyTorch fallback exists: a non-CUDA tensor raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import IOU_INFER, IOU_METRICS, ORDER_CLASS_GROUPED, ORDER_GLOBAL_CONF  # noqa: F401

S_, B_, C_ = 7, 2, 20


def _check_pred(t: torch.Tensor, name: str, S: int, B: int, Cn: int) -> torch.Tensor:
    _lib.require_cuda(t, name)
    D = B * 5 + Cn
    if t.dim() == 3:
        t = t.unsqueeze(0)
    if t.dim() != 4 or tuple(t.shape[1:]) != (S, S, D):
        raise ValueError(f"{name}: expected shape (N,{S},{S},{D}), got {tuple(t.shape)}")
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    if t.data_ptr() % 8:
        t = t.clone()
    return t


def _thr_array(thresholds):
    thr = [float(t) for t in thresholds]
    if not 1 <= len(thr) <= _lib.MAX_THRESHOLDS:
        raise NotImplementedError(
            f"{len(thr)} IoU thresholds: the matching kernel carries 1..{_lib.MAX_THRESHOLDS} "
            "(4 ground-truth subsets x T + 1 flag bits per detection)")
    return thr, (C.c_double * len(thr))(*thr)


@dataclass
class DecodeResult:
    n_keep: torch.Tensor      # (N,) int32
    keep_k: torch.Tensor      # (N,K) uint8 candidate indices, ordered
    keep_cls: torch.Tensor    # (N,K) uint8
    keep_conf: torch.Tensor   # (N,K) float64
    keep_box: torch.Tensor    # (N,K,4) float64 (x_abs, y_abs, w, h)
    n_cand: torch.Tensor      # (N,) int32 candidates above conf_thr
    bad_range: torch.Tensor   # (N,) uint8


def decode_nms(pred: torch.Tensor, conf_thr: float, nms_thr: float, iou_mode: int = IOU_METRICS,
               order: int = ORDER_CLASS_GROUPED, S: int = S_, B: int = B_, Cn: int = C_,
               want_boxes: bool = True) -> DecodeResult:
    """yh_decode_nms over a batch (or a single (S,S,D) image)."""
    p = _check_pred(pred, "predictions", S, B, Cn)
    N, K, dev = p.shape[0], S * S * B, p.device
    n_keep = torch.zeros(N, dtype=torch.int32, device=dev)
    n_cand = torch.zeros(N, dtype=torch.int32, device=dev)
    bad = torch.zeros(N, dtype=torch.uint8, device=dev)
    keep_k = torch.empty(N, K, dtype=torch.uint8, device=dev)
    keep_cls = torch.empty(N, K, dtype=torch.uint8, device=dev)
    keep_conf = torch.empty(N, K, dtype=torch.float64, device=dev)
    keep_box = torch.empty(N, K, 4, dtype=torch.float64, device=dev) if want_boxes else None
    with _lib.on_device(dev):
        rc = _lib.lib().yh_decode_nms(p.data_ptr(), N, S, B, Cn, float(conf_thr), float(nms_thr), int(iou_mode),
                                      int(order), n_keep.data_ptr(), keep_k.data_ptr(), keep_cls.data_ptr(),
                                      keep_conf.data_ptr(), _lib.ptr(keep_box), n_cand.data_ptr(),
                                      bad.data_ptr(), _lib.stream_ptr(dev))
    _lib.check(rc, "yh_decode_nms")
    return DecodeResult(n_keep, keep_k, keep_cls, keep_conf, keep_box, n_cand, bad)


@dataclass
class UpdateChunk:
    """Device-resident result of one mAPMetric.update() batch."""
    n: int
    n_keep: torch.Tensor     # (N,) int32
    rec_key: torch.Tensor    # (N,K) uint64-as-int64: ~orderable(conf)
    rec_flags: torch.Tensor  # (N,K) int64: class << 56 | TP bits
    rec_k: torch.Tensor      # (N,K) uint8 candidate index
    rec_raw: torch.Tensor | None  # (N,K,3) int64 words: x_abs, y_abs (fp64 bits), (w, h) fp32 pair
    n_gt: torch.Tensor       # (N,) int32
    gt_cell: torch.Tensor | None  # (N,S*S) uint8
    gt_cls: torch.Tensor | None   # (N,S*S) uint8
    gt_raw: torch.Tensor | None   # (N,S*S,4) float32
    shape: tuple = (S_, B_, C_)
    # without list data: the batch itself (device tensors + their version counters at update time), from which
    # the reference's list form can be rebuilt on demand
    src: tuple | None = None


def map_update(pred: torch.Tensor, tgt: torch.Tensor, conf_thr: float, nms_thr: float, thresholds,
               gt_count: torch.Tensor, S: int = S_, B: int = B_, Cn: int = C_,
               want_lists: bool = False) -> UpdateChunk:
    """yh_map_update: fused decode + NMS + GT parse + TP/FP flags; adds into gt_count (4*C int32).
    want_lists: also write the optional list-form side data (24-byte boxes, ground-truth cells / classes /
    raw boxes) that yh_list_form expands; without it the chunk keeps a reference to the batch instead."""
    p = _check_pred(pred, "predictions", S, B, Cn)
    t = _check_pred(tgt, "targets", S, B, Cn)
    if p.shape != t.shape:
        raise ValueError(f"predictions {tuple(p.shape)} and targets {tuple(t.shape)} differ in shape")
    _, thr_c = _thr_array(thresholds)
    N, K, G, dev = p.shape[0], S * S * B, S * S, p.device
    i64 = dict(dtype=torch.int64, device=dev)
    u8 = dict(dtype=torch.uint8, device=dev)
    n_keep = torch.zeros(N, dtype=torch.int32, device=dev)
    n_gt = torch.zeros(N, dtype=torch.int32, device=dev)
    rec_key = torch.empty(N, K, **i64)
    rec_flags = torch.empty(N, K, **i64)
    rec_k = torch.empty(N, K, **u8)
    rec_raw = torch.empty(N, K, 3, dtype=torch.int64, device=dev) if want_lists else None
    gt_cell = torch.empty(N, G, **u8) if want_lists else None
    gt_cls = torch.empty(N, G, **u8) if want_lists else None
    gt_raw = torch.empty(N, G, 4, dtype=torch.float32, device=dev) if want_lists else None
    with _lib.on_device(dev):
        rc = _lib.lib().yh_map_update(p.data_ptr(), t.data_ptr(), N, S, B, Cn, float(conf_thr), float(nms_thr),
                                      thr_c, len(thr_c), n_keep.data_ptr(), rec_key.data_ptr(),
                                      rec_flags.data_ptr(), rec_k.data_ptr(), _lib.ptr(rec_raw), n_gt.data_ptr(),
                                      _lib.ptr(gt_cell), _lib.ptr(gt_cls), _lib.ptr(gt_raw), gt_count.data_ptr(),
                                      _lib.stream_ptr(dev))
    _lib.check(rc, "yh_map_update")
    src = None if want_lists else (p, t, p._version, t._version)
    return UpdateChunk(N, n_keep, rec_key, rec_flags, rec_k, rec_raw, n_gt, gt_cell, gt_cls, gt_raw, (S, B, Cn), src)


_FUSED_WS = _lib.Workspace(zeroed=True)


def map_update_loss(pred: torch.Tensor, tgt: torch.Tensor, conf_thr: float, nms_thr: float, thresholds,
                    gt_count: torch.Tensor, lambda_coord: float, lambda_noobj: float, want_grad: bool = True,
                    tail_stream: "torch.cuda.Stream | None" = None, phases: int = 3):
    """yh_map_update_loss: the update kernel with the loss (forward + backward) riding on it; one read of the two
    tensors.  S=7, B=2, C=20 and fp32 predictions.  Returns (chunk, parts (5,) fp32 device tensor, grad or None).
    tail_stream: run the loss tail (object cells, fold) there, forked behind the update kernel, so that whatever the
    caller enqueues next on the current stream (compaction, exchange, sort) starts right away; `parts` and the object
    cells of `grad` are then ready on tail_stream, which the caller joins.  phases (measurement only): 1 = the update
    kernel alone, 2 = the loss tail alone (on what the last phase-1 call left in the workspace), 3 = both."""
    p = _check_pred(pred, "predictions", S_, B_, C_)
    t = _check_pred(tgt, "targets", S_, B_, C_)
    if p.shape != t.shape:
        raise ValueError(f"predictions {tuple(p.shape)} and targets {tuple(t.shape)} differ in shape")
    if p.shape[0] == 0:
        raise ValueError("empty batch")
    _, thr_c = _thr_array(thresholds)
    N, K, dev = p.shape[0], S_ * S_ * B_, p.device
    n_keep = torch.zeros(N, dtype=torch.int32, device=dev)
    n_gt = torch.zeros(N, dtype=torch.int32, device=dev)
    rec_key = torch.empty(N, K, dtype=torch.int64, device=dev)
    rec_flags = torch.empty(N, K, dtype=torch.int64, device=dev)
    rec_k = torch.empty(N, K, dtype=torch.uint8, device=dev)
    parts = torch.empty(5, dtype=torch.float32, device=dev)
    grad = torch.empty_like(p) if want_grad else None
    lib = _lib.lib()
    ws = _FUSED_WS.get(dev, int(lib.yh_map_update_loss_workspace_bytes(N)))

    def call(phases, stream_handle):
        rc = lib.yh_map_update_loss(p.data_ptr(), t.data_ptr(), N, S_, B_, C_, float(conf_thr), float(nms_thr), thr_c,
                                    len(thr_c), n_keep.data_ptr(), rec_key.data_ptr(), rec_flags.data_ptr(),
                                    rec_k.data_ptr(), n_gt.data_ptr(), gt_count.data_ptr(), float(lambda_coord),
                                    float(lambda_noobj), parts.data_ptr(), _lib.ptr(grad), ws.data_ptr(), ws.numel(),
                                    phases, stream_handle)
        _lib.check(rc, "yh_map_update_loss")

    with _lib.on_device(dev):
        if tail_stream is None:
            call(int(phases), _lib.stream_ptr(dev))
        else:
            call(1, _lib.stream_ptr(dev))
            tail_stream.wait_stream(torch.cuda.current_stream(dev))
            call(2, tail_stream.cuda_stream)
    chunk = UpdateChunk(N, n_keep, rec_key, rec_flags, rec_k, None, n_gt, None, None, None, (S_, B_, C_),
                        (p, t, p._version, t._version))
    return chunk, parts, grad


def parse_gt(tgt: torch.Tensor, S: int = S_, B: int = B_, Cn: int = C_):
    """yh_parse_gt over a batch: (n_gt int32 (N,), gt_cls uint8 (N,S*S), gt_box float64 (N,S*S,4))."""
    t = _check_pred(tgt, "targets", S, B, Cn)
    N, G, dev = t.shape[0], S * S, t.device
    n_gt = torch.zeros(N, dtype=torch.int32, device=dev)
    gt_cls = torch.empty(N, G, dtype=torch.uint8, device=dev)
    gt_box = torch.empty(N, G, 4, dtype=torch.float64, device=dev)
    with _lib.on_device(dev):
        rc = _lib.lib().yh_parse_gt(t.data_ptr(), N, S, B, Cn, n_gt.data_ptr(), gt_cls.data_ptr(), gt_box.data_ptr(),
                                    _lib.stream_ptr(dev))
    _lib.check(rc, "yh_parse_gt")
    return n_gt, gt_cls, gt_box


def list_form(ch: UpdateChunk):
    """yh_list_form: (rec_pos uint8 (N,K), rec_box float64 (N,K,4), gt_box float64 (N,S*S,4))."""
    if ch.rec_raw is None:
        raise _lib.YoloHeadError("this batch was accumulated without list data (want_lists=False)")
    S, B, Cn = ch.shape
    N, K = ch.rec_key.shape
    dev = ch.rec_key.device
    rec_pos = torch.empty(N, K, dtype=torch.uint8, device=dev)
    rec_box = torch.empty(N, K, 4, dtype=torch.float64, device=dev)
    gt_box = torch.empty(N, S * S, 4, dtype=torch.float64, device=dev)
    with _lib.on_device(dev):
        rc = _lib.lib().yh_list_form(ch.n_keep.data_ptr(), ch.rec_key.data_ptr(), ch.rec_flags.data_ptr(),
                                     ch.rec_k.data_ptr(), ch.rec_raw.data_ptr(), ch.n_gt.data_ptr(),
                                     ch.gt_cell.data_ptr(), ch.gt_raw.data_ptr(), N, S, B, Cn, rec_pos.data_ptr(),
                                     rec_box.data_ptr(), gt_box.data_ptr(), _lib.stream_ptr(dev))
    _lib.check(rc, "yh_list_form")
    return rec_pos, rec_box, gt_box


def compact_records(chunk: UpdateChunk, out_key: torch.Tensor | None = None, out_flags: torch.Tensor | None = None,
                    running: torch.Tensor | None = None):
    """Dense (key, flags) arrays of a chunk in image order.

    Stand-alone: returns (key, flags, offsets) with capacity N*K and offsets[N] = number of
    records (a device scalar).  Chained: pass shared out_key / out_flags buffers and `running`
    (device int64 scalar = records already there); this batch is appended and `running` advanced
    on the stream, without a host round trip."""
    N, K = chunk.rec_key.shape
    dev = chunk.rec_key.device
    lib = _lib.lib()
    offsets = torch.empty(N + 1, dtype=torch.int64, device=dev)
    if out_key is None:
        out_key = torch.empty(N * K, dtype=torch.int64, device=dev)
        out_flags = torch.empty(N * K, dtype=torch.int64, device=dev)
    ws = _COMPACT_WS.get(dev, lib.yh_compact_workspace_bytes(N))
    with _lib.on_device(dev):
        rc = lib.yh_compact_records(chunk.n_keep.data_ptr(), chunk.rec_key.data_ptr(), chunk.rec_flags.data_ptr(),
                                    N, K, offsets.data_ptr(), out_key.data_ptr(), out_flags.data_ptr(),
                                    _lib.ptr(running), ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "yh_compact_records")
    return out_key, out_flags, offsets


_AP_WS = _lib.Workspace()
_COMPACT_WS = _lib.Workspace()


def ap_reduce(key: torch.Tensor, flags: torch.Tensor, M: int, gt_count: torch.Tensor, Cn: int, T: int,
              m_dev: torch.Tensor | None = None):
    """yh_ap_reduce over the first M records (or min(*m_dev, M) when the count lives on the
    device).  key/flags are clobbered.

    Returns (ap (4,C,T) float64, totals (2,) int64, packed float64 [4*C*T + 3 + 4*C]) on the device;
    `packed` holds ap, totals, an error flag and gt_count, so one copy brings everything to the host."""
    dev = gt_count.device
    lib = _lib.lib()
    nbytes = lib.yh_ap_workspace_bytes(int(M), Cn, T)
    if nbytes == 0:
        raise NotImplementedError(f"yh_ap_reduce: C={Cn}, T={T} outside this build")
    ws = _AP_WS.get(dev, nbytes)
    ap = torch.empty(4, Cn, T, dtype=torch.float64, device=dev)
    totals = torch.empty(2, dtype=torch.int64, device=dev)
    packed = torch.empty(4 * Cn * T + 3 + 4 * Cn, dtype=torch.float64, device=dev)
    with _lib.on_device(dev):
        rc = lib.yh_ap_reduce(key.data_ptr() if M else None, flags.data_ptr() if M else None, int(M),
                              _lib.ptr(m_dev), gt_count.data_ptr(), Cn, T, ap.data_ptr(), totals.data_ptr(),
                              packed.data_ptr(), ws.data_ptr(), nbytes, _lib.stream_ptr(dev))
    _lib.check(rc, "yh_ap_reduce")
    return ap, totals, packed


def unpack_ap(packed: torch.Tensor, Cn: int, T: int):
    """Host view of ap_reduce's packed buffer: (ap ndarray (4,C,T), tp_total, n_records, gt_count ndarray (4,C))."""
    h = packed.cpu().numpy()
    n = 4 * Cn * T
    if h[n + 2] != 0.0:
        raise _lib.YoloHeadError("yh_ap_reduce: the chained scan of the radix sort gave up (internal error)")
    return h[:n].reshape(4, Cn, T), int(h[n]), int(h[n + 1]), h[n + 3:].astype("int64").reshape(4, Cn)


def iou_pairs(box_a, box_b, iou_mode: int, device) -> torch.Tensor:
    a = torch.as_tensor(box_a, dtype=torch.float64).reshape(-1, 4).to(device)
    b = torch.as_tensor(box_b, dtype=torch.float64).reshape(-1, 4).to(device)
    _lib.require_cuda(a, "boxes")
    out = torch.empty(a.shape[0], dtype=torch.float64, device=a.device)
    with _lib.on_device(a.device):
        rc = _lib.lib().yh_iou_pairs(a.data_ptr(), b.data_ptr(), a.shape[0], int(iou_mode), out.data_ptr(),
                                     _lib.stream_ptr(a.device))
    _lib.check(rc, "yh_iou_pairs")
    return out


def nms_lists(offsets, cls, conf, box, nms_thr: float, iou_mode: int, order: int, device):
    """Greedy NMS over CSR lists (host sequences in, device tensors out: n_keep, keep_idx)."""
    off = torch.as_tensor(offsets, dtype=torch.int64)
    n_lists = off.numel() - 1
    max_len = int((off[1:] - off[:-1]).max()) if n_lists > 0 else 0
    off_d = off.to(device)
    _lib.require_cuda(off_d, "lists")
    cls_d = torch.as_tensor(cls, dtype=torch.int32).to(device)
    conf_d = torch.as_tensor(conf, dtype=torch.float64).to(device)
    box_d = torch.as_tensor(box, dtype=torch.float64).reshape(-1, 4).to(device)
    n_keep = torch.zeros(max(n_lists, 1), dtype=torch.int32, device=device)
    keep_idx = torch.zeros(max(cls_d.numel(), 1), dtype=torch.int32, device=device)
    with _lib.on_device(off_d.device):
        rc = _lib.lib().yh_nms_lists(off_d.data_ptr(), n_lists, cls_d.data_ptr(), conf_d.data_ptr(),
                                     box_d.data_ptr(), float(nms_thr), int(iou_mode), int(order),
                                     n_keep.data_ptr(), keep_idx.data_ptr(), max_len,
                                     _lib.stream_ptr(off_d.device))
    _lib.check(rc, "yh_nms_lists")
    return n_keep[:n_lists], keep_idx


def match_lists(det_off, det_cls, det_conf, det_box, gt_off, gt_cls, gt_box, Cn: int, thresholds,
                gt_count: torch.Tensor):
    """yh_match_lists; all list arguments are device tensors.  Returns (rec_key, rec_flags)."""
    _, thr_c = _thr_array(thresholds)
    dev = gt_count.device
    n_lists = det_off.numel() - 1
    M = det_cls.numel()
    max_dets = int((det_off[1:] - det_off[:-1]).max()) if n_lists > 0 else 0
    rec_key = torch.empty(max(M, 1), dtype=torch.int64, device=dev)
    rec_flags = torch.empty(max(M, 1), dtype=torch.int64, device=dev)
    with _lib.on_device(dev):
        rc = _lib.lib().yh_match_lists(det_off.data_ptr(), det_cls.data_ptr(), det_conf.data_ptr(),
                                       det_box.data_ptr(), gt_off.data_ptr(), gt_cls.data_ptr(),
                                       gt_box.data_ptr(), n_lists, Cn, thr_c, len(thr_c), rec_key.data_ptr(),
                                       rec_flags.data_ptr(), gt_count.data_ptr(), max_dets,
                                       _lib.stream_ptr(dev))
    _lib.check(rc, "yh_match_lists")
    return rec_key, rec_flags
