"""BASELINE.json configs[3] in miniature: YOLOLoss inside a real training step.

The reference trains with `loss, loss_dict = criterion(predictions, targets); loss.backward()`
(training/trainer.py:72-90, optionally under autocast + GradScaler).  These tests run that
step on a small conv + detection-head model with the fused loss kernel and with the torch
restatement of the reference loss (oracle/loss_oracle.py) from identical weights, and compare
what training actually consumes: parameter gradients and the weights after optimizer steps.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

pytestmark = pytest.mark.gpu


class TinyYolo(nn.Module):
    """Conv trunk + the reference's head shape (Flatten -> Linear -> LeakyReLU -> Linear -> view,
    models.py:326-348), small enough to run in milliseconds."""

    def __init__(self, S=7, B=2, C=20):
        super().__init__()
        self.S, self.B, self.C = S, B, C
        self.trunk = nn.Sequential(nn.Conv2d(3, 8, 3, stride=2, padding=1), nn.LeakyReLU(0.1),
                                   nn.Conv2d(8, 16, 3, stride=2, padding=1), nn.LeakyReLU(0.1),
                                   nn.AdaptiveAvgPool2d(S))
        self.head = nn.Sequential(nn.Flatten(), nn.Linear(16 * S * S, 128), nn.LeakyReLU(0.1),
                                  nn.Linear(128, S * S * (B * 5 + C)))

    def forward(self, x):
        return self.head(self.trunk(x)).view(-1, self.S, self.S, self.B * 5 + self.C)


def _data(n, seed):
    from yolo_v1_b200 import synth

    g = torch.Generator().manual_seed(seed)
    images = torch.randn(n, 3, 56, 56, generator=g)
    targets = synth.make_targets(n, seed=seed, p_obj=0.08, both_slots=0.05)
    return images, targets


def _fresh_model(seed=0):
    torch.manual_seed(seed)
    return TinyYolo()


def _oracle_total(pred, tgt):
    from oracle import loss_oracle

    return loss_oracle.yolo_loss(pred, tgt)[0]


def _train(loss_kind, steps, images, targets, dev, amp=None):
    """`steps` SGD steps from the same initial weights; returns (losses, final parameters, first-step grads)."""
    from yolo_v1_b200 import YOLOLoss

    model = _fresh_model().to(dev)
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
    crit = YOLOLoss()
    scaler = torch.amp.GradScaler("cuda", init_scale=1024.0) if amp == torch.float16 else None
    losses, first = [], None
    x, t = images.to(dev), targets.to(dev)
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=amp, enabled=amp is not None):
            pred = model(x)
            if loss_kind == "kernel":
                loss, parts = crit(pred, t)
                assert set(parts) == {"total", "coord", "conf_obj", "conf_noobj", "class"}
            else:
                loss = _oracle_total(pred.float(), t)
        if scaler is not None:
            scaler.scale(loss).backward()
            scaler.unscale_(opt)
        else:
            loss.backward()
        if first is None:
            first = [p.grad.detach().clone() for p in model.parameters()]
        if scaler is not None:
            scaler.step(opt)
            scaler.update()
        else:
            opt.step()
        losses.append(float(loss.detach()))
    return losses, [p.detach().clone() for p in model.parameters()], first


def _close(a, b, rtol, atol_frac=1e-6):
    scale = float(b.abs().max()) or 1.0
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol_frac * scale)


def test_parameter_gradients_and_sgd_steps_match_the_reference_loss():
    dev = torch.device("cuda")
    images, targets = _data(48, seed=11)
    lk, wk, gk = _train("kernel", 5, images, targets, dev)
    lo, wo, go = _train("oracle", 5, images, targets, dev)
    # cuDNN's weight-gradient reductions are not bit-reproducible run to run: tolerances are
    # relative to each tensor's largest entry, an order above that noise
    for a, b in zip(gk, go):          # what backward hands the optimizer
        _close(a, b, rtol=1e-4, atol_frac=1e-4)
    for a, b in zip(wk, wo):          # weights after five momentum-SGD steps
        _close(a, b, rtol=1e-5, atol_frac=1e-5)
    assert lk == pytest.approx(lo, rel=1e-5)
    assert lk[-1] < lk[0]             # and it trains


@pytest.mark.parametrize("amp", [torch.bfloat16, torch.float16])
def test_autocast_training_step(amp):
    """Half-precision predictions (trainer.py's autocast path): the kernel reads them as they are,
    returns an fp32 loss and a gradient in the prediction dtype; GradScaler's scaled backward works."""
    dev = torch.device("cuda")
    images, targets = _data(32, seed=5)
    lk, wk, gk = _train("kernel", 3, images, targets, dev, amp=amp)
    lo, wo, go = _train("oracle", 3, images, targets, dev, amp=amp)
    tol = 3e-2 if amp == torch.bfloat16 else 5e-3      # the two graphs round differently in 8 / 11 bits
    for a, b in zip(gk, go):
        _close(a, b, rtol=tol, atol_frac=tol)
    assert lk == pytest.approx(lo, rel=tol)
    assert all(torch.isfinite(w).all() for w in wk)


# ----------------------------------------------------------------------------- DDP
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _ddp_worker(rank, world, port, backend, q):
    from torch.nn.parallel import DistributedDataParallel as DDP

    from yolo_v1_b200 import YOLOLoss

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    kw = {"device_id": dev} if backend == "nccl" else {}
    dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    try:
        images, targets = _data(32, seed=3)
        per = 32 // world
        x, t = images[rank * per:(rank + 1) * per].to(dev), targets[rank * per:(rank + 1) * per].to(dev)
        model = DDP(_fresh_model().to(dev), device_ids=[dev.index] if backend == "nccl" else None)
        loss, _ = YOLOLoss()(model(x), t)
        loss.backward()                       # DDP all-reduces (averages) the parameter gradients
        # numpy arrays travel by value (tensors would be shared through file descriptors of a process about to exit)
        q.put((rank, [p.grad.detach().cpu().numpy() for p in model.parameters()], float(loss.detach())))
    finally:
        dist.destroy_process_group()


def test_ddp_gradients_equal_the_full_batch_gradients():
    """Image-sharded training (configs[3]): each rank's loss is a mean over ITS images, DDP averages
    the gradients, which is the gradient of the full-batch mean when the shards are equal."""
    two_gpus = torch.cuda.device_count() >= 2
    world, backend = 2, ("nccl" if two_gpus else "gloo")   # one GPU: both ranks share it, gloo carries the grads
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ddp_worker, args=(r, world, port, backend, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {r: (g, l) for r, g, l in (q.get(timeout=300) for _ in range(world))}
    for p in procs:
        p.join(60)
    from yolo_v1_b200 import YOLOLoss

    dev = torch.device("cuda")
    images, targets = _data(32, seed=3)
    model = _fresh_model().to(dev)
    loss, _ = YOLOLoss()(model(images.to(dev)), targets.to(dev))
    loss.backward()
    full = [p.grad.detach().cpu() for p in model.parameters()]
    assert (got[0][1] + got[1][1]) / 2 == pytest.approx(float(loss.detach()), rel=1e-5)
    for r in range(world):
        for a, b in zip(got[r][0], full):     # 16 + 16 images vs 32 at once: another summation order
            _close(torch.from_numpy(a), b, rtol=1e-4, atol_frac=1e-4)
