"""Torch fp32 restatement of the YOLOv1 loss (CPU checker for the CUDA loss kernel).

TEST INFRASTRUCTURE ONLY -- never imported by yolo_v1_b200/.

Follows /root/reference/src/yolo/loss.py:
  object cells / target slot selection   loss.py:98-104
  IoU of each predicted box vs target    loss.py:107-111, 174-212 (fp32, +1e-6 in the denominator)
  responsible box = first arg-max IoU    loss.py:110-118
  coord / conf_obj / conf_noobj / class  loss.py:126-159
  total = sum / N, loss_dict already /N  loss.py:162-170
The gradient is whatever autograd derives from this graph, which is the
reference's own definition of backward (it has no hand-written one); the IoU
target is NOT detached (loss.py:111,123,144).

Written with dense masks (torch.where) instead of the reference's boolean
indexing, so fp32 summation order differs in the last bits; pinned against the
reference's outputs in tests/test_oracle_golden.py at rtol 1e-6.
"""
from __future__ import annotations

import torch


def iou_xywh(p: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    """p (...,B,4), t (...,1,4) centre-format boxes -> IoU (...,B); loss.py:191-212."""
    px1, px2 = p[..., 0] - p[..., 2] / 2, p[..., 0] + p[..., 2] / 2
    py1, py2 = p[..., 1] - p[..., 3] / 2, p[..., 1] + p[..., 3] / 2
    tx1, tx2 = t[..., 0] - t[..., 2] / 2, t[..., 0] + t[..., 2] / 2
    ty1, ty2 = t[..., 1] - t[..., 3] / 2, t[..., 1] + t[..., 3] / 2
    iw = torch.clamp(torch.minimum(px2, tx2) - torch.maximum(px1, tx1), min=0)
    ih = torch.clamp(torch.minimum(py2, ty2) - torch.maximum(py1, ty1), min=0)
    inter = iw * ih
    union = p[..., 2] * p[..., 3] + t[..., 2] * t[..., 3] - inter
    return inter / (union + 1e-6)


def yolo_loss(pred: torch.Tensor, tgt: torch.Tensor, S=7, B=2, C=20, lambda_coord=5.0,
              lambda_noobj=0.5):
    """Returns (total 0-dim tensor attached to pred's graph, parts tensor[5] detached).

    parts = (total, coord, conf_obj, conf_noobj, class), each already divided by N.
    """
    N = pred.shape[0]
    pb = pred[..., : B * 5].reshape(N, S, S, B, 5)
    tb = tgt[..., : B * 5].reshape(N, S, S, B, 5)
    pc, tc = pred[..., B * 5:], tgt[..., B * 5:]

    filled = tb[..., 4] > 0                       # (N,S,S,B)
    obj = filled.any(dim=-1)                      # (N,S,S)
    slot = filled.float().argmax(dim=-1)          # first filled slot, 0 if none
    tbox = torch.gather(tb[..., :4], 3, slot[..., None, None].expand(N, S, S, 1, 4)).squeeze(3)

    ious = iou_xywh(pb[..., :4], tbox.unsqueeze(3))           # (N,S,S,B)
    best = ious.argmax(dim=-1)
    best_iou = ious.gather(-1, best.unsqueeze(-1)).squeeze(-1)
    resp = torch.zeros_like(ious, dtype=torch.bool).scatter_(-1, best.unsqueeze(-1), True)
    resp = resp & obj.unsqueeze(-1)

    zero = torch.zeros((), dtype=ious.dtype, device=pred.device)
    d_xy = (pb[..., :2] - tbox.unsqueeze(3)[..., :2]) ** 2
    d_wh = (torch.sqrt(torch.clamp(pb[..., 2:4], min=1e-6))
            - torch.sqrt(torch.clamp(tbox.unsqueeze(3)[..., 2:4], min=1e-6))) ** 2
    coord = lambda_coord * (torch.where(resp.unsqueeze(-1), d_xy, zero).sum()
                            + torch.where(resp.unsqueeze(-1), d_wh, zero).sum())
    conf_obj = torch.where(resp, (pb[..., 4] - best_iou.unsqueeze(-1)) ** 2, zero).sum()
    conf_noobj = lambda_noobj * torch.where(~resp, pb[..., 4] ** 2, zero).sum()
    cls = torch.where(obj.unsqueeze(-1), (pc - tc) ** 2, zero).sum()

    total = (coord + conf_obj + conf_noobj + cls) / N
    parts = torch.stack([total, coord / N, conf_obj / N, conf_noobj / N, cls / N]).detach()
    return total, parts


def yolo_loss_and_grad(pred: torch.Tensor, tgt: torch.Tensor, **kw):
    """Convenience: (parts[5] float32, grad wrt pred in pred's dtype)."""
    p = pred.detach().clone().requires_grad_(True)
    total, parts = yolo_loss(p, tgt, **kw)
    total.backward()
    return parts, p.grad.detach()
