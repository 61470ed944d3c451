"""CapturedHeadStep: the whole head step as one CUDA graph gives the eager step's results, bit for bit,
also after new data is written into the captured tensors, and across ranks with the peer exchange."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _eager(pred, tgt, conf):
    from yolo_v1_b200 import YOLOLoss, mAPMetric

    p = pred.detach().clone().requires_grad_(True)
    total, parts = YOLOLoss()(p, tgt)
    total.backward()
    m = mAPMetric(num_classes=20, conf_threshold=conf)
    m.update(p.detach(), tgt)
    return parts, {k: float(v) for k, v in m.compute().items()}, p.grad


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("profile,conf", [("correlated", 0.01), ("uniform", 0.3)])
def test_replay_equals_eager_and_follows_new_data(profile, conf, fused):
    from yolo_v1_b200 import CapturedHeadStep, YOLOLoss, mAPMetric, synth

    dev = torch.device("cuda")
    n = 700
    sets = []
    for seed in (1, 2, 3):
        t = synth.make_targets(n, seed=seed, p_obj=0.055, size_mix=True)
        sets.append((synth.make_preds(n, profile, seed=seed, targets=t).to(dev), t.to(dev)))
    pred, tgt = sets[0][0].clone(), sets[0][1].clone()
    step = CapturedHeadStep(YOLOLoss(), mAPMetric(num_classes=20, conf_threshold=conf), pred, tgt, fused=fused)
    assert step.launches_per_replay >= 10 and step.fused == fused
    for p_new, t_new in sets + sets[:1]:
        pred.copy_(p_new)
        tgt.copy_(t_new)
        loss, res = step.replay()
        want_loss, want_res, want_grad = _eager(p_new, t_new, conf)
        if fused:
            # the loss rides on the update kernel: the same per-cell arithmetic (the gradient is bit-identical),
            # the five sums are added up in a different order than the stand-alone loss kernel's
            assert loss.keys() == want_loss.keys()
            for k in loss:
                assert loss[k] == pytest.approx(want_loss[k], rel=2e-6, abs=1e-9), k
        else:
            assert loss == want_loss
        assert {k: float(v) for k, v in res.items()} == want_res
        assert torch.equal(step.grad, want_grad)


def test_pipelined_submit_collect_returns_every_step_in_order():
    """Two alternating graphs: step k + 1 is queued before step k is decoded.  New data is written into the captured
    tensors on the stream between submissions; every step must come back, in order, equal to the eager step."""
    from yolo_v1_b200 import CapturedHeadStep, YOLOLoss, _lib, mAPMetric, synth

    dev = torch.device("cuda")
    n = 600
    sets = []
    for seed in (11, 12, 13, 14, 15):
        t = synth.make_targets(n, seed=seed, p_obj=0.055, size_mix=True)
        sets.append((synth.make_preds(n, "correlated", seed=seed, targets=t).to(dev), t.to(dev)))
    pred, tgt = sets[0][0].clone(), sets[0][1].clone()
    step = CapturedHeadStep(YOLOLoss(), mAPMetric(num_classes=20, conf_threshold=0.01), pred, tgt, pipelined=True)
    want = [_eager(p, t, 0.01) for p, t in sets]
    got, grads = [], []
    pred.copy_(sets[0][0]); tgt.copy_(sets[0][1])
    step.submit()
    for i in range(len(sets)):
        if i + 1 < len(sets):
            pred.copy_(sets[i + 1][0]); tgt.copy_(sets[i + 1][1])      # stream-ordered behind step i
            step.submit()
        loss, res = step.collect()
        got.append((loss, {k: float(v) for k, v in res.items()}))
        grads.append(step.grad.clone())
    for (loss, res), g, (w_loss, w_res, w_grad) in zip(got, grads, want):
        assert res == w_res and torch.equal(g, w_grad)
        assert all(loss[k] == pytest.approx(w_loss[k], rel=2e-6, abs=1e-9) for k in w_loss)
    with pytest.raises(_lib.YoloHeadError):
        step.collect()                      # nothing in flight
    step.submit(); step.submit()
    with pytest.raises(_lib.YoloHeadError):
        step.submit()                       # both slots in flight
    step.collect(); step.collect()
    loss, res = step.replay()               # the strict form still works on a pipelined object
    assert {k: float(v) for k, v in res.items()} == got[-1][1]


def test_forward_only_capture():
    from yolo_v1_b200 import CapturedHeadStep, YOLOLoss, mAPMetric, synth

    dev = torch.device("cuda")
    t = synth.make_targets(300, seed=9, p_obj=0.055, size_mix=True)
    p = synth.make_preds(300, "correlated", seed=9, targets=t).to(dev)
    for fused in (True, False):
        step = CapturedHeadStep(YOLOLoss(), mAPMetric(num_classes=20), p, t.to(dev), backward=False, fused=fused)
        loss, res = step.replay()
        want_loss, want_res, _ = _eager(p, t.to(dev), 0.01)
        assert all(loss[k] == pytest.approx(want_loss[k], rel=2e-6, abs=1e-9) for k in want_loss)
        assert fused or loss == want_loss
        assert {k: float(v) for k, v in res.items()} == want_res and step.grad is None


# ----------------------------------------------------------------------------- two ranks, peer exchange in the graph
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, backend, q):
    from yolo_v1_b200 import CapturedHeadStep, YOLOLoss, mAPMetric, synth
    from yolo_v1_b200 import distributed as D

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    kw = {"device_id": dev} if backend == "nccl" else {}
    dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    try:
        n = 640
        lo, hi = D.shard_range(n, rank, world)
        data = []
        for seed in (4, 5):
            t = synth.make_targets(n, seed=seed, p_obj=0.055, size_mix=True)
            data.append((synth.make_preds(n, "correlated", seed=seed, targets=t)[lo:hi].to(dev), t[lo:hi].to(dev)))
        pred, tgt = data[0][0].clone(), data[0][1].clone()
        dm = D.DistributedMAP(mAPMetric(num_classes=20, conf_threshold=0.01), mode="auto")
        step = CapturedHeadStep(YOLOLoss(), dm, pred, tgt)
        out = []
        for p_new, t_new in data + data[:1]:
            pred.copy_(p_new)
            tgt.copy_(t_new)
            loss, res = step.replay()
            out.append({k: float(v) for k, v in res.items()})
        q.put((rank, out, step.launches_per_replay, dm.mode))
        dm.close()
    finally:
        dist.destroy_process_group()


def test_peer_exchange_inside_the_graph():
    from yolo_v1_b200 import mAPMetric, synth

    two = torch.cuda.device_count() >= 2
    world, backend = 2, ("nccl" if two else "gloo")       # one GPU: the two ranks time-slice it
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, backend, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {r: (o, l, m) for r, o, l, m in (q.get(timeout=300) for _ in range(world))}
    for p in procs:
        p.join(60)
    want = []
    for seed in (4, 5):
        t = synth.make_targets(640, seed=seed, p_obj=0.055, size_mix=True)
        p = synth.make_preds(640, "correlated", seed=seed, targets=t)
        m = mAPMetric(num_classes=20, conf_threshold=0.01)
        m.update(p.cuda(), t.cuda())
        want.append({k: float(v) for k, v in m.compute().items()})
    want.append(want[0])
    for r in range(world):
        out, launches, mode = got[r]
        assert mode == "peer" and launches >= 15
        assert out == want
