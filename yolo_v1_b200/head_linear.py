"""The detection head's last Linear fused with the loss: host side of yh_head_linear_loss.

The reference ends `DetectionHead.forward` with `nn.Linear(4096, S*S*(B*5+C))` and a `view`
(src/yolo/models.py:326-331, 345-348) and hands the result to `YOLOLoss` (src/yolo/loss.py:55-172); under
`autocast("cuda")` (src/yolo/training/trainer.py:70, 165) the layer runs on fp16 operands.  Here the two are one
tcgen05 kernel: the loss (and its gradient) is evaluated on the accumulator tile, so the [M, 1470] prediction tensor
only goes to memory when the caller wants it (`want_pred`, for the metric).

    fused = FusedHeadLoss(head.fc_layers[-1], criterion)            # the nn.Linear itself: shares its parameters
    loss, loss_dict = fused(features_4096, targets)                  # as criterion(head(...), targets)
    loss.backward()                                                  # features / weight / bias gradients

`head_linear(x, weight, bias)` is the plain layer (fp32 [M, 7, 7, 30] out), for the inference path.
There is no fallback: other shapes or dtypes raise NotImplementedError and the caller keeps its two calls.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from .loss import _KEYS, LossDict, YOLOLoss

S, B, C, IN_FEATURES = 7, 2, 20, 4096
TF32 = 3   # YH_TF32
OUT_FEATURES = S * S * (B * 5 + C)

_WS = _lib.Workspace(zeroed=True)


def _check_operands(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]):
    _lib.require_cuda(x, "x")
    _lib.require_cuda(weight, "weight")
    if x.dim() != 2 or x.shape[1] != IN_FEATURES or tuple(weight.shape) != (OUT_FEATURES, IN_FEATURES):
        raise NotImplementedError(f"head_linear: built for Linear({IN_FEATURES} -> {OUT_FEATURES}); got x {tuple(x.shape)}, "
                                  f"weight {tuple(weight.shape)}")
    if x.dtype not in (torch.float16, torch.bfloat16, torch.float32) or weight.dtype is not x.dtype:
        raise NotImplementedError(f"head_linear: fp16, bf16 or fp32 operands of one type (x {x.dtype}, weight {weight.dtype}); "
                                  "cast them as autocast would")
    if bias is not None and (bias.dtype is not torch.float32 or bias.numel() != OUT_FEATURES):
        raise _lib.YoloHeadError("head_linear: bias must be fp32 [1470]")
    return x.contiguous(), weight.contiguous(), None if bias is None else bias.contiguous()


def linear_loss(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], targets: Optional[torch.Tensor],
                lambda_coord: float = 5.0, lambda_noobj: float = 0.5, grad_scale: float = 1.0, round_pred: bool = True,
                want_grad: bool = True, grad_dtype: Optional[torch.dtype] = None, want_pred: bool = False,
                precision: str = "3xtf32"):
    """One launch of yh_head_linear_loss.  Returns (parts, grad, pred): parts (5,) fp32 device tensor in loss_dict order,
    grad [M, 1470] = grad_scale * d total / d prediction (None without want_grad), pred fp32 [M, 7, 7, 30] (None without
    want_pred).  targets=None: the plain layer (pred only).  fp32 operands (the reference's default, non-AMP step): precision
    "3xtf32" = three tensor-core passes over the operands' TF32 and residual parts, fp32's accuracy; "tf32" = one pass
    (what torch.backends.cuda.matmul.allow_tf32 gives); the gradient is fp32 and round_pred does not apply."""
    x, weight, bias = _check_operands(x, weight, bias)
    dev, M = x.device, x.shape[0]
    if targets is not None:
        _lib.require_cuda(targets, "targets")
        if targets.dtype is not torch.float32 or tuple(targets.shape) != (M, S, S, B * 5 + C):
            raise _lib.YoloHeadError(f"linear_loss: targets must be fp32 [{M}, 7, 7, 30], got {targets.dtype} {tuple(targets.shape)}")
        targets = targets.contiguous()
    else:
        want_grad, want_pred = False, True
    gdt = x.dtype if grad_dtype is None else grad_dtype
    if x.dtype is torch.float32:
        if precision not in ("3xtf32", "tf32"):
            raise ValueError(f"linear_loss: precision must be '3xtf32' or 'tf32', got {precision!r}")
        if gdt is not torch.float32:
            raise _lib.YoloHeadError("linear_loss: fp32 operands take an fp32 gradient")
    code = _lib.dtype_code(x) if x.dtype is not torch.float32 else (_lib.F32 if precision == "3xtf32" else TF32)
    if M == 0:   # sums over nothing divided by a batch size of 0 (loss.py:160-162): nan, and nothing to launch
        return (None if targets is None else torch.full((5,), float("nan"), device=dev),
                torch.empty(0, OUT_FEATURES, dtype=gdt, device=dev) if want_grad else None,
                torch.empty(0, S, S, B * 5 + C, device=dev) if want_pred else None)
    with _lib.on_device(dev):
        parts = torch.empty(5, dtype=torch.float32, device=dev) if targets is not None else None
        grad = torch.empty(M, OUT_FEATURES, dtype=gdt, device=dev) if want_grad else None
        pred = torch.empty(M, S, S, B * 5 + C, dtype=torch.float32, device=dev) if want_pred else None
        lib = _lib.lib()
        nbytes = lib.yh_head_linear_loss_workspace_bytes()
        ws = _WS.get(dev, nbytes)
        rc = lib.yh_head_linear_loss(x.data_ptr(), weight.data_ptr(), _lib.ptr(bias), code, _lib.ptr(targets), M,
                                     S, B, C, IN_FEATURES, lambda_coord, lambda_noobj, grad_scale, int(bool(round_pred)),
                                     _lib.ptr(parts), _lib.ptr(grad), _lib._DTYPES[gdt], _lib.ptr(pred), ws.data_ptr(), nbytes,
                                     _lib.stream_ptr(dev))
        _lib.check(rc, "yh_head_linear_loss")
    return parts, grad, pred


def head_linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, round_pred: bool = True,
                precision: str = "3xtf32") -> torch.Tensor:
    """`linear(x).view(-1, 7, 7, 30)` (models.py:326-331, 345-348) as fp32, ready for YOLOInference / mAPMetric."""
    return linear_loss(x, weight, bias, None, round_pred=round_pred, precision=precision)[2]


class _FusedFn(torch.autograd.Function):
    """Inputs are the caller's tensors (fp32 master weights, say); the cast to the operand type happens in here, so the
    gradients go back in the callers' own types and an fp32 backward is not rounded to the operand type on the way."""

    @staticmethod
    def forward(ctx, x, weight, bias, targets, operand_dtype, lambda_coord, lambda_noobj, grad_scale, round_pred,
                backward_dtype, want_pred, precision):
        need = x.requires_grad or weight.requires_grad or (bias is not None and bias.requires_grad)
        xo, wo = x.detach().to(operand_dtype), weight.detach().to(operand_dtype)
        bo = None if bias is None else bias.detach().to(torch.float32)
        parts, grad, pred = linear_loss(xo, wo, bo, targets, lambda_coord, lambda_noobj, grad_scale, round_pred,
                                        want_grad=need, grad_dtype=backward_dtype, want_pred=want_pred, precision=precision)
        ctx.save_for_backward(xo, wo, grad)
        ctx.grad_scale = grad_scale
        # what `grad` has been multiplied by since the kernel wrote it (a device scalar: no host read in backward)
        ctx.applied = torch.ones(1, dtype=torch.float32, device=xo.device) if grad is not None else None
        ctx.in_dtypes = (x.dtype, weight.dtype, None if bias is None else bias.dtype)
        ctx.mark_non_differentiable(parts)
        if pred is not None:
            ctx.mark_non_differentiable(pred)
        return parts[0].clone(), parts, pred

    @staticmethod
    def backward(ctx, g_total, _g_parts, _g_pred):
        xo, wo, grad = ctx.saved_tensors
        if grad is None:
            return (None,) * 12
        # grad already carries grad_scale; what arrives here is the scaler's factor again (or 1): bring the buffer to
        # upstream / grad_scale IN PLACE (yh_scale_inplace returns at once when the factor is exactly 1, the usual case;
        # a second backward over a retained graph finds what the first one applied in ctx.applied)
        want = (g_total.detach().to(torch.float32) / ctx.grad_scale).reshape(1)
        factor = (want / ctx.applied).contiguous()
        with _lib.on_device(grad.device):
            rc = _lib.lib().yh_scale_inplace(grad.data_ptr(), _lib.dtype_code(grad), grad.numel(), factor.data_ptr(),
                                             _lib.stream_ptr(grad.device))
        _lib.check(rc, "yh_scale_inplace")
        ctx.applied.copy_(want)
        g = grad
        dx, dw, db = ctx.in_dtypes
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            gx = (g @ wo.to(g.dtype)).to(dx)
        if ctx.needs_input_grad[1]:
            gw = (g.t() @ xo.to(g.dtype)).to(dw)
        if db is not None and ctx.needs_input_grad[2]:
            gb = g.to(torch.float32).sum(0).to(db)
        return (gx, gw, gb) + (None,) * 9


class FusedHeadLoss(torch.nn.Module):
    """`criterion(linear(x).view(-1, S, S, B*5+C), targets)` in one kernel.

    linear: the head's last nn.Linear(4096, 1470) (its parameters are used in place: fp32 master weights are cast to
    `operand_dtype` per call, as autocast does; an fp16 / bf16 layer is used as is).  criterion: a YOLOLoss (S=7, B=2,
    C=20).  backward_dtype: the type d loss / d prediction is written in and the two backward GEMMs run in -- None = the
    operand type (autocast's behaviour), torch.float32 = fp32 GEMMs.  operand_dtype=torch.float32 is the reference's default
    (non-AMP) step: fp32 operands through three TF32 passes (`precision`, see linear_loss), fp32 backward GEMMs.  loss_scale: the GradScaler factor in effect
    (`scaler.get_scale()`), so that an fp16 gradient does not underflow; the backward divides it out of what
    `scaler.scale(loss).backward()` sends."""

    def __init__(self, linear: torch.nn.Linear, criterion: YOLOLoss, operand_dtype: torch.dtype = torch.float16,
                 round_pred: bool = True, backward_dtype: Optional[torch.dtype] = None, precision: str = "3xtf32"):
        super().__init__()
        if (linear.in_features, linear.out_features) != (IN_FEATURES, OUT_FEATURES) or (criterion.S, criterion.B, criterion.C) != (S, B, C):
            raise NotImplementedError("FusedHeadLoss: built for Linear(4096 -> 1470) and YOLOLoss(S=7, B=2, C=20)")
        self.linear, self.criterion = linear, criterion
        self.operand_dtype, self.round_pred, self.backward_dtype = operand_dtype, round_pred, backward_dtype
        self.precision = precision
        self.last_pred: Optional[torch.Tensor] = None

    def forward(self, x: torch.Tensor, targets: torch.Tensor, loss_scale: float = 1.0,
                want_pred: bool = False) -> Tuple[torch.Tensor, Dict[str, float]]:
        w = self.linear.weight
        dt = w.dtype if w.dtype in (torch.float16, torch.bfloat16) else self.operand_dtype
        b = self.linear.bias
        total, parts, pred = _FusedFn.apply(x, w, b, targets, dt, self.criterion.lambda_coord, self.criterion.lambda_noobj,
                                            float(loss_scale), self.round_pred, self.backward_dtype, want_pred, self.precision)
        self.last_pred = pred
        if self.criterion.defer_loss_dict:
            return total, LossDict(parts)
        return total, dict(zip(_KEYS, parts.tolist()))   # the one device->host read the loss_dict API requires
