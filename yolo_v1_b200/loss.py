"""YOLOLoss -- drop-in for the reference's loss module, computed by one fused CUDA kernel.

Same constructor, attributes, return value and loss_dict keys as
/root/reference/src/yolo/loss.py:7-172 (``YOLOLoss(S, B, C, lambda_coord,
lambda_noobj)``; ``forward(predictions, targets) -> (total, dict)``;
``compute_iou`` static method, loss.py:174-212).  forward + backward are a single
pass of ``yh_loss_fwd_bwd`` (csrc/loss.cu); the autograd backward only multiplies
the saved gradient by the upstream scalar (``yh_scale_inplace``, which exits
immediately when that scalar is 1).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib

_KEYS = ("total", "coord", "conf_obj", "conf_noobj", "class")


_WS = _lib.Workspace(zeroed=True)


def loss_fwd_bwd(pred: torch.Tensor, tgt: torch.Tensor, S: int, B: int, C: int, lambda_coord: float,
                 lambda_noobj: float, want_grad: bool = True):
    """Raw kernel call: returns (parts[5] fp32 device tensor, grad or None)."""
    _lib.require_cuda(pred, "predictions")
    _lib.require_cuda(tgt, "targets")
    D = B * 5 + C
    if pred.dim() != 4 or tuple(pred.shape[1:]) != (S, S, D) or pred.shape != tgt.shape:
        raise ValueError(f"expected predictions and targets of shape (N,{S},{S},{D}); got "
                         f"{tuple(pred.shape)} and {tuple(tgt.shape)}")
    N = pred.shape[0]
    if N == 0:
        raise ValueError("empty batch")
    p = pred.detach()
    if not p.is_contiguous():
        p = p.contiguous()
    t = tgt.detach()
    if t.dtype is not torch.float32 or not t.is_contiguous():
        t = t.to(torch.float32).contiguous()
    lib = _lib.lib()
    parts = torch.empty(5, dtype=torch.float32, device=p.device)
    grad = torch.empty_like(p) if want_grad else None
    nbytes = lib.yh_loss_workspace_bytes(N, S, B, C)
    if nbytes == 0:
        raise NotImplementedError(f"YOLOLoss: shape S={S}, B={B}, C={C} is outside this build (B <= 8)")
    ws = _WS.get(p.device, nbytes)
    with _lib.on_device(p.device):
        rc = lib.yh_loss_fwd_bwd(p.data_ptr(), _lib.dtype_code(p), t.data_ptr(), N, S, B, C,
                                 float(lambda_coord), float(lambda_noobj), parts.data_ptr(), _lib.ptr(grad),
                                 ws.data_ptr(), ws.numel(), _lib.stream_ptr(p.device))
    _lib.check(rc, "yh_loss_fwd_bwd")
    return parts, grad


class LossDict(dict):
    """The loss_dict of YOLOLoss.forward: a dict[str, float] with the reference's five keys
    (loss.py:164-170) whose values are copied from the device the first time anything is read.

    The reference calls .item() five times inside forward, which stalls the host on every
    training step; here the kernel's five floats stay on the device until somebody looks, so
    forward + backward are enqueued without a host round trip.  Reading any entry (or iterating,
    printing, comparing, copying, pickling) does the one 20-byte copy and from then on this is a
    plain dict.  Opt-in (YOLOLoss(defer_loss_dict=True)): C-level fast paths that look inside a
    dict without calling its methods (json's C encoder, for one) would see it empty before the
    first read, so the default stays the reference's eager dict of floats."""

    __slots__ = ("_parts",)

    def __init__(self, parts: torch.Tensor):
        super().__init__()
        self._parts = parts

    def _fill(self):
        parts = self._parts
        if parts is not None:
            self._parts = None
            dict.update(self, zip(_KEYS, parts.tolist()))

    def _wrap(name):  # noqa: N805 - helper run at class-creation time
        base = getattr(dict, name)

        def method(self, *a, **kw):
            self._fill()
            return base(self, *a, **kw)

        method.__name__ = name
        return method

    for _n in ("__getitem__", "__iter__", "__len__", "__contains__", "__repr__", "__eq__", "__ne__", "__or__",
               "__ror__", "__setitem__", "__delitem__", "__reversed__", "keys", "values", "items", "get", "pop",
               "popitem", "setdefault", "update", "copy", "clear"):
        locals()[_n] = _wrap(_n)
    del _n, _wrap

    def __reduce__(self):
        self._fill()
        return (dict, (dict(self),))


class _YoloLossFn(torch.autograd.Function):
    """forward runs the fused kernel once (loss parts + d loss / d predictions); backward hands that
    gradient to autograd scaled by the upstream scalar.  The buffer is given away on the first
    backward (so autograd can adopt it without a copy); a further backward over a retained graph
    (`retain_graph=True`) re-runs the kernel on the saved inputs, as the reference's graph would
    recompute its gradient."""

    @staticmethod
    def forward(ctx, pred, tgt, S, B, C, lc, ln):
        need = pred.requires_grad
        parts, grad = loss_fwd_bwd(pred, tgt, S, B, C, lc, ln, want_grad=need)
        ctx.saved_grad = grad
        ctx.cfg = (S, B, C, lc, ln)
        if need:
            ctx.save_for_backward(pred, tgt)    # references only; autograd checks their versions
        ctx.mark_non_differentiable(parts)
        return parts[0].clone(), parts

    @staticmethod
    def backward(ctx, g_total, _g_parts):
        if not ctx.needs_input_grad[0]:
            return (None,) * 7
        grad = ctx.saved_grad
        ctx.saved_grad = None
        if grad is None:                        # second backward: the first one gave the buffer away
            pred, tgt = ctx.saved_tensors
            grad = loss_fwd_bwd(pred, tgt, *ctx.cfg, want_grad=True)[1]
        scale = g_total.detach().to(torch.float32).reshape(1).contiguous()
        with _lib.on_device(grad.device):
            rc = _lib.lib().yh_scale_inplace(grad.data_ptr(), _lib.dtype_code(grad), grad.numel(),
                                             scale.data_ptr(), _lib.stream_ptr(grad.device))
        _lib.check(rc, "yh_scale_inplace")
        return (grad, None, None, None, None, None, None)


class YOLOLoss(nn.Module):
    """YOLO v1 loss (reference: src/yolo/loss.py:7-53 for the constructor contract)."""

    def __init__(self, S: int = 7, B: int = 2, C: int = 20, lambda_coord: float = 5.0,
                 lambda_noobj: float = 0.5, *, defer_loss_dict: bool = False):
        super().__init__()
        self.S, self.B, self.C = S, B, C
        self.lambda_coord = lambda_coord
        self.lambda_noobj = lambda_noobj
        # extension (keyword-only): return a LossDict that is read from the device when first used
        self.defer_loss_dict = defer_loss_dict

    def forward(self, predictions: torch.Tensor, targets: torch.Tensor) -> tuple[torch.Tensor, dict[str, float]]:
        total, parts = _YoloLossFn.apply(predictions, targets, self.S, self.B, self.C, self.lambda_coord,
                                         self.lambda_noobj)
        if self.defer_loss_dict:
            return total, LossDict(parts)
        vals = parts.tolist()  # the one device->host read the loss_dict API requires (loss.py:164-170)
        return total, dict(zip(_KEYS, vals))

    @staticmethod
    def compute_iou(boxes1: torch.Tensor, boxes2: torch.Tensor) -> torch.Tensor:
        """IoU of (…,B,4) vs (…,1,4) centre-format boxes, +1e-6 in the denominator
        (loss.py:174-212).  Public helper of the reference; the fused kernel computes the
        same quantity internally, this stays as plain tensor ops for API parity."""
        b1x1, b1x2 = boxes1[..., 0] - boxes1[..., 2] / 2, boxes1[..., 0] + boxes1[..., 2] / 2
        b1y1, b1y2 = boxes1[..., 1] - boxes1[..., 3] / 2, boxes1[..., 1] + boxes1[..., 3] / 2
        b2x1, b2x2 = boxes2[..., 0] - boxes2[..., 2] / 2, boxes2[..., 0] + boxes2[..., 2] / 2
        b2y1, b2y2 = boxes2[..., 1] - boxes2[..., 3] / 2, boxes2[..., 1] + boxes2[..., 3] / 2
        iw = torch.clamp(torch.minimum(b1x2, b2x2) - torch.maximum(b1x1, b2x1), min=0)
        ih = torch.clamp(torch.minimum(b1y2, b2y2) - torch.maximum(b1y1, b2y1), min=0)
        inter = iw * ih
        union = boxes1[..., 2] * boxes1[..., 3] + boxes2[..., 2] * boxes2[..., 3] - inter
        return inter / (union + 1e-6)
