"""Multi-rank exchange logic of yolo_v1_b200.distributed.

CPU part (gloo, world_size 2 and 3): sharding arithmetic and both record exchanges,
checked against a single-process concatenation.  GPU part: the sharded metric on
2 ranks against the single-GPU metric and the CPU oracle, bit for bit.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from yolo_v1_b200 import distributed as D  # noqa: E402


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_ranges_cover_the_set_contiguously():
    for n in (0, 1, 7, 64, 4952, 65536):
        for world in (1, 2, 3, 4, 8):
            spans = [D.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 == b0 and a0 <= a1 and b0 <= b1
    assert D.shard_range(4952, 7, 8) == (4333, 4952)


def test_class_owners_are_contiguous_blocks():
    for world in (1, 2, 3, 4, 8, 32):
        own = D.class_owners(20, world)
        assert len(own) == 20 and own == sorted(own) and max(own) < world
    own = D.class_owners(20, 8)          # balanced: nobody idle, nobody above ceil(20 / 8)
    assert own == [0, 0, 0, 1, 1, 2, 2, 2, 3, 3, 4, 4, 4, 5, 5, 6, 6, 6, 7, 7]
    assert sorted(own.count(r) for r in range(8)) == [2, 2, 2, 2, 3, 3, 3, 3]


def _records(rank, n):
    """Synthetic compact records of one rank: key = confidence key, flags = class << 56 | serial."""
    rng = np.random.default_rng(100 + rank)
    key = rng.integers(0, 8, size=n, dtype=np.int64)             # few distinct keys -> many ties
    cls = rng.integers(0, 20, size=n, dtype=np.int64)
    serial = (np.arange(n, dtype=np.int64) + (rank << 24))       # where the record came from
    return torch.from_numpy(key), torch.from_numpy((cls << 56) | serial)


def _worker(rank, world, port, counts, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        key, flg = _records(rank, counts[rank])
        # ---- mode "gather"
        gk, gf, cnt = D.gather_records(key, flg)
        assert cnt == list(counts)
        exp_k = torch.cat([_records(r, counts[r])[0] for r in range(world)])
        exp_f = torch.cat([_records(r, counts[r])[1] for r in range(world)])
        assert torch.equal(gk, exp_k) and torch.equal(gf, exp_f)
        # ---- mode "class_shard": numpy stands in for yh_partition_records (a stable partition)
        owners = np.array(D.class_owners(20, world))
        own = owners[(flg.numpy() >> 56) & 0xFF]
        perm = np.argsort(own, kind="stable")
        part = [int((own == p).sum()) for p in range(world)]
        mk, mf = D.exchange_by_owner(key[perm], flg[perm], part)
        sel = owners[(exp_f.numpy() >> 56) & 0xFF] == rank        # my classes, global image order
        assert torch.equal(mk, exp_k[sel]) and torch.equal(mf, exp_f[sel])
        # ---- counts / reductions used by compute()
        t = torch.tensor([counts[rank]], dtype=torch.int64)
        dist.all_reduce(t)
        assert int(t) == sum(counts)
        out.put((rank, "ok"))
    except Exception as e:  # pragma: no cover - surfaced in the parent
        out.put((rank, repr(e)))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("counts", [(1000, 777), (0, 50), (300, 0, 411)])
def test_record_exchanges_over_gloo(counts):
    world = len(counts)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, counts, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
    res = sorted(q.get(timeout=5) for _ in range(world))
    assert res == [(r, "ok") for r in range(world)], res
    assert all(p.exitcode == 0 for p in procs)


# ----------------------------------------------------------------------------- GPU
def _gpu_worker(rank, world, port, n, mode, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank % torch.cuda.device_count())
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from yolo_v1_b200 import mAPMetric, synth

        tgt = synth.make_targets(n, seed=7, p_obj=0.055, size_mix=True)
        pred = synth.make_preds(n, "correlated", seed=7, targets=tgt)
        res = D.evaluate_sharded(None, pred, tgt, batch=97, mode=mode, num_classes=20, conf_threshold=0.01)
        q.put((rank, {k: float(v) for k, v in res.items()}))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["gather", "class_shard", "peer"])
def test_sharded_metric_equals_single_gpu_and_oracle(mode):
    from oracle import c_oracle, results
    from yolo_v1_b200 import mAPMetric, synth

    n = 500
    # 2 ranks; on a 1-GPU box both ranks share cuda:0 (NCCL refuses that), so fall back to world 1 there
    world = 2 if torch.cuda.device_count() >= 2 else 1
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gpu_worker, args=(r, world, port, n, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(60)
    tgt = synth.make_targets(n, seed=7, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(n, "correlated", seed=7, targets=tgt)
    single = mAPMetric(num_classes=20, conf_threshold=0.01)
    single.update(pred.cuda(), tgt.cuda())
    want = {k: float(v) for k, v in single.compute().items()}
    st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), 0.01, 0.4)
    ref = c_oracle.compute(st, single.iou_thresholds)
    orc = results.assemble(ref["ap"], ref["overall"], ref["sizes"], single.iou_thresholds, 20, n)
    for r in range(world):
        assert got[r].keys() == want.keys()
        for k in want:
            assert got[r][k] == want[k], (mode, r, k, got[r][k], want[k])     # bit-exact vs one GPU
            assert abs(got[r][k] - float(orc[k])) <= 1e-12 * max(1.0, abs(float(orc[k]))), (k, got[r][k], orc[k])


def _peer_growth_worker(rank, world, port, backend, sizes, q, image_exchange=True, late_rank=-1):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    kw = {"device_id": dev} if backend == "nccl" else {}
    dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    try:
        from yolo_v1_b200 import mAPMetric, synth

        dm = D.DistributedMAP(mAPMetric(num_classes=20, conf_threshold=0.01), mode="peer")
        dm.image_exchange = image_exchange
        out = []
        for n in sizes:
            tgt = synth.make_targets(n, seed=n, p_obj=0.055, size_mix=True)
            pred = synth.make_preds(n, "correlated", seed=n, targets=tgt)
            lo, hi = D.shard_range(n, rank, world)
            dm.reset()
            for s in range(lo, hi, 128):
                dm.update(pred[s:min(hi, s + 128)].to(dev), tgt[s:min(hi, s + 128)].to(dev))
            if rank == late_rank:
                import time

                time.sleep(5.5)      # longer than the 4 s a flag wait lasts
            res = dm.compute()
            out.append(({k: float(v) for k, v in res.items()}, dm.peer_fallbacks, dm._peer.capacity))
        dm.close()
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


def _run_peer_growth(world, backend, sizes, image_exchange=True, late_rank=-1):
    from yolo_v1_b200 import mAPMetric, synth

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_growth_worker, args=(r, world, port, backend, sizes, q, image_exchange, late_rank)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(60)
    want = []
    for n in sizes:
        tgt = synth.make_targets(n, seed=n, p_obj=0.055, size_mix=True)
        pred = synth.make_preds(n, "correlated", seed=n, targets=tgt)
        single = mAPMetric(num_classes=20, conf_threshold=0.01)
        single.update(pred.cuda(), tgt.cuda())
        want.append({k: float(v) for k, v in single.compute().items()})
    return got, want


@pytest.mark.gpu
def test_peer_blocks_overflow_falls_back_and_grows():
    """The receive blocks are sized by the first compute(); a later, larger set overflows them:
    that call takes the NCCL exchange (same bits) and the blocks are re-created larger."""
    world = 2 if torch.cuda.device_count() >= 2 else 1
    got, want = _run_peer_growth(world, "nccl", (40, 600, 600))   # 600 does not fit blocks sized for 40
    for r in range(world):
        (r0, f0, c0), (r1, f1, c1), (r2, f2, c2) = got[r]
        assert (f0, f1, f2) == (0, 1, 1) and c1 > c0 and c2 == c1, (f0, f1, f2, c0, c1, c2)
        for res, w in zip((r0, r1, r2), want):
            assert res == w


@pytest.mark.gpu
@pytest.mark.parametrize("world,image_exchange", [(2, True), (2, False), (3, True), (5, True)])
def test_peer_exchange_between_processes_sharing_one_gpu(world, image_exchange):
    """CUDA IPC mapping + the flag protocol with several ranks even on a one-GPU box (gloo carries
    the handles; the kernels of the processes time-slice the device).  image_exchange: records go from
    the update batches straight to their owners (yh_exchange_images) -- or compacted first
    (yh_compact_records + yh_exchange_records); both must equal the single-process result bit for bit."""
    got, want = _run_peer_growth(world, "gloo", (600, 333, 40, 7), image_exchange)
    for r in range(world):
        assert [f for _, f, _ in got[r]] == [0, 0, 0, 0]
        for (res, _, _), w in zip(got[r], want):
            assert res == w


@pytest.mark.gpu
def test_a_late_rank_does_not_time_the_peer_exchange_out():
    """One rank reaches compute() 5.5 s after the other; the exchange's flag waits last 4 s.  The eager
    compute() orders the ranks on the stream first, so the early rank waits there (as it would in NCCL)."""
    got, want = _run_peer_growth(2, "gloo", (300,), True, late_rank=1)
    for r in range(2):
        assert got[r][0][0] == want[0] and got[r][0][1] == 0
