"""Golden vectors for the fused head Linear + loss, from the UNMODIFIED reference (/root/reference/src).

    python tests/golden/make_head_linear_golden.py        ->  tests/golden/head_linear.npz

What the reference computes at the end of DetectionHead.forward and in the criterion (models.py:326-331, 345-348;
loss.py:55-172), on CPU in fp32 over operands that are exactly representable in bf16 AND fp16 (8 significant bits), so
that the same inputs can be fed to the GPU kernel in either type:
    pred = nn.Linear(4096, 1470)(x).view(-1, 7, 7, 30);  loss, loss_dict = YOLOLoss()(pred, targets);  loss.backward()
Inputs are regenerated from a seed by yolo_v1_b200.synth.exact_uniform (same bits on any machine); stored are the loss
dict, the bias gradient, slices of the prediction / activation / weight gradients, and the same loss under
torch.autocast(bfloat16) (the layer's output rounded to bf16 before the loss).  Build container only.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/src")

from yolo.loss import YOLOLoss  # noqa: E402  (reference)

from yolo_v1_b200 import synth  # noqa: E402

KEYS = ("total", "coord", "conf_obj", "conf_noobj", "class")


inputs = synth.head_linear_inputs


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    out = {}
    cases = [("m64", 64, 5), ("m200", 200, 6), ("m333", 333, 7)]
    for name, M, seed in cases:
        x, w, b, tgt = inputs(M, seed)
        for t in (torch.bfloat16, torch.float16):
            assert torch.equal(torch.from_numpy(x).to(t).float(), torch.from_numpy(x)), "x not exact"
            assert torch.equal(torch.from_numpy(w).to(t).float(), torch.from_numpy(w)), "w not exact"
        lin = torch.nn.Linear(4096, 1470)
        with torch.no_grad():
            lin.weight.copy_(torch.from_numpy(w))
            lin.bias.copy_(torch.from_numpy(b))
        xt = torch.from_numpy(x).requires_grad_(True)
        tg = torch.from_numpy(tgt)
        crit = YOLOLoss()
        pred = lin(xt).view(-1, 7, 7, 30)
        pred.retain_grad()
        loss, d = crit(pred, tg)
        loss.backward()
        out[f"{name}/M"] = np.int64(M)
        out[f"{name}/seed"] = np.int64(seed)
        out[f"{name}/parts"] = np.array([d[k] for k in KEYS], dtype=np.float64)
        out[f"{name}/pred_head"] = pred.detach()[:4].numpy()
        out[f"{name}/dpred_head"] = pred.grad[:4].numpy()
        out[f"{name}/dpred_abs_sum"] = np.float64(pred.grad.double().abs().sum().item())
        out[f"{name}/db"] = lin.bias.grad.numpy()
        out[f"{name}/dx_head"] = xt.grad[:8, :256].numpy()
        out[f"{name}/dx_abs_sum"] = np.float64(xt.grad.double().abs().sum().item())
        out[f"{name}/dw_head"] = lin.weight.grad[:64, :128].numpy()
        out[f"{name}/dw_abs_sum"] = np.float64(lin.weight.grad.double().abs().sum().item())
        # the autocast flavour: the layer's output is bf16 when the criterion sees it
        with torch.autocast("cpu", dtype=torch.bfloat16):
            pred_a = lin(torch.from_numpy(x)).view(-1, 7, 7, 30)
        assert pred_a.dtype is torch.bfloat16
        _, da = crit(pred_a, tg)
        out[f"{name}/parts_autocast_bf16"] = np.array([da[k] for k in KEYS], dtype=np.float64)
        # and the reference criterion on the bf16-rounded prediction widened back to fp32 (what round_pred computes)
        _, dr = crit(pred.detach().to(torch.bfloat16).float(), tg)
        out[f"{name}/parts_rounded_bf16"] = np.array([dr[k] for k in KEYS], dtype=np.float64)
        print(name, {k: round(d[k], 6) for k in KEYS}, "autocast", round(da["total"], 5), "rounded", round(dr["total"], 5))
    out["cases"] = np.array([c[0] for c in cases])
    np.savez_compressed(os.path.join(HERE, "head_linear.npz"), **out)


if __name__ == "__main__":
    main()
