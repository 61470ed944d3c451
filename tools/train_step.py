"""BASELINE.json configs[3]: ResNet50-backbone YOLOv1 training step, random init, 448x448, 32 images per GPU.

    python tools/train_step.py [--amp bf16|fp16|none] [--batch 32] [--steps 10]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_step.py   (DDP, NCCL)

The backbone and the detection head stay stock PyTorch / cuDNN (north-star); only the loss is
this repo's kernel.  Reported: full step (forward, loss, backward, gradient all-reduce under DDP,
SGD step) in ms and images/s; the same step with the loss swapped for a trivial mean-square of the
head output (what the model alone costs); the loss kernel alone on the head output of this batch.
Architecture as the reference builds it (models.py:128-176, 281-348): torchvision resnet50 without
avgpool/fc -> 4 convs (2048->1024, one stride 2) with LeakyReLU(0.1) -> Flatten -> Linear(50176, 4096)
-> LeakyReLU -> Dropout(0.5) -> Linear(4096, 1470) -> view(N, 7, 7, 30).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import YOLOLoss, synth  # noqa: E402


class ResNetYolo(nn.Module):
    def __init__(self, S=7, B=2, C=20):
        super().__init__()
        from torchvision.models import resnet50

        self.S, self.B, self.C = S, B, C
        self.backbone = nn.Sequential(*list(resnet50(weights=None).children())[:-2])
        act = lambda: nn.LeakyReLU(0.1)  # noqa: E731
        self.convs = nn.Sequential(nn.Conv2d(2048, 1024, 3, padding=1), act(),
                                   nn.Conv2d(1024, 1024, 3, stride=2, padding=1), act(),
                                   nn.Conv2d(1024, 1024, 3, padding=1), act(),
                                   nn.Conv2d(1024, 1024, 3, padding=1), act())
        self.fc = nn.Sequential(nn.Flatten(), nn.Linear(1024 * S * S, 4096), act(), nn.Dropout(0.5),
                                nn.Linear(4096, S * S * (B * 5 + C)))

    def forward(self, x, targets=None, fused=None, loss_scale=1.0):
        if fused is None:
            return self.fc(self.convs(self.backbone(x))).view(-1, self.S, self.S, self.B * 5 + self.C)
        # --fused-head: everything up to the last Linear here, the layer + view + loss in yolo_v1_b200.FusedHeadLoss
        # (inside forward so that DDP sees the layer's parameters being used)
        return fused(self.fc[:-1](self.convs(self.backbone(x))), targets, loss_scale=loss_scale)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--amp", default="bf16", choices=["none", "bf16", "fp16"])
    ap.add_argument("--batch", type=int, default=32, help="images per GPU")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--fused-head", action="store_true",
                    help="the head's last Linear + view + loss as one kernel (yolo_v1_b200.FusedHeadLoss) instead of Linear, then YOLOLoss")
    a = ap.parse_args()
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1234)
    model = ResNetYolo().to(dev).to(memory_format=torch.channels_last)
    if world > 1:
        model = nn.parallel.DistributedDataParallel(model, device_ids=[local])
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, weight_decay=5e-4)
    crit = YOLOLoss(defer_loss_dict=True)
    amp = {"none": None, "bf16": torch.bfloat16, "fp16": torch.float16}[a.amp]
    fused = None
    if a.fused_head:
        from yolo_v1_b200 import FusedHeadLoss
        inner = model.module if world > 1 else model
        fused = FusedHeadLoss(inner.fc[-1], crit, operand_dtype=amp if amp is not None else torch.float32)
    scaler = torch.amp.GradScaler("cuda") if amp == torch.float16 else None
    g = torch.Generator().manual_seed(1234 + rank)
    x = torch.randn(a.batch, 3, 448, 448, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
    t = synth.make_targets(a.batch, seed=1234 + rank, p_obj=0.055, size_mix=True).to(dev)

    import contextlib

    def step(use_kernel, sync=True):
        opt.zero_grad(set_to_none=True)
        ctx = model.no_sync() if (world > 1 and not sync) else contextlib.nullcontext()
        with ctx:
            return _step(use_kernel)

    def _step(use_kernel):
        with torch.autocast("cuda", dtype=amp, enabled=amp is not None):
            if use_kernel and fused is not None:
                loss, parts = model(x, t, fused, scaler.get_scale() if scaler is not None else 1.0)
                pred = None
            else:
                pred = model(x)
                if use_kernel:
                    loss, parts = crit(pred, t)
                else:
                    loss, parts = pred.float().square().mean(), None
        if scaler is not None:
            scaler.scale(loss).backward()
            scaler.step(opt)
            scaler.update()
        else:
            loss.backward()
            opt.step()
        return loss, parts, pred

    def timed(fn, k):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / k
        if world > 1:
            m = torch.tensor([ms], device=dev)
            dist.all_reduce(m, op=dist.ReduceOp.MAX)
            ms = float(m)
        return ms, out

    for _ in range(a.warmup):
        step(True)
        step(False)
    full_ms, (loss, parts, pred) = timed(lambda: step(True), a.steps)
    bare_ms, _ = timed(lambda: step(False), a.steps)
    nosync_ms = None
    if world > 1:   # the same step without DDP's gradient all-reduce: the difference is what NCCL adds to the step
        step(True, sync=False)
        nosync_ms, _ = timed(lambda: step(True, sync=False), a.steps)
    if pred is None:   # --fused-head: the loss-alone timing below still wants a head output
        with torch.no_grad(), torch.autocast("cuda", dtype=amp, enabled=amp is not None):
            pred = model(x)
    head_out = pred.detach()

    def loss_only():
        p = head_out.clone().requires_grad_(True)
        l, _ = crit(p, t)
        l.backward()
    clone_ms, _ = timed(lambda: head_out.clone(), 50)
    lk_ms, _ = timed(loss_only, 50)
    if rank == 0:
        print(json.dumps({
            "fused_head": bool(a.fused_head),
            "config": f"ResNet50-YOLOv1 training step, random init, {a.batch} x 448x448 per GPU, amp={a.amp}, "
                      f"{world} GPU(s){', DDP/NCCL' if world > 1 else ''}, SGD momentum",
            "step_ms": full_ms, "images_per_s": world * a.batch / (full_ms * 1e-3),
            "step_ms_with_trivial_loss": bare_ms, "loss_share_of_step_ms": full_ms - bare_ms,
            "loss_fwd_bwd_alone_ms": lk_ms - clone_ms, "head_output_dtype": str(head_out.dtype),
            "loss_total": float(parts["total"]), "n_gpus": world, "global_batch": world * a.batch,
            "step_ms_without_gradient_allreduce": nosync_ms,
            "exposed_allreduce_ms": (full_ms - nosync_ms) if nosync_ms is not None else None,
            "gradient_bytes": sum(p.numel() * p.element_size() for p in model.parameters()),
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
