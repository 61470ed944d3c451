"""Adversarial inputs for the fused update kernel: flags / keeps vs the CPU oracle, bit-exact.

Dense same-class scenes (49 ground truths and 98 detections of one class), NaN and
infinite values, every threshold count the flag word carries, unaligned views, and
full-size (65,536 images) invariants that need no oracle."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, results
from yolo_v1_b200 import _lib, head, mAPMetric, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"
DEFAULT_THR = [0.5 + 0.05 * i for i in range(10)]


def _check(pred, tgt, conf_thr, thr=DEFAULT_THR, nms=0.4):
    n = pred.shape[0]
    st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), conf_thr, nms, n_threads=4)
    ref = c_oracle.compute(st, thr, n_threads=4)
    m = mAPMetric(num_classes=20, iou_thresholds=list(thr), conf_threshold=conf_thr, nms_threshold=nms)
    m.update(pred.to(DEV), tgt.to(DEV))
    ch = m._chunks[0]
    n_keep = ch.n_keep.cpu().numpy()
    np.testing.assert_array_equal(n_keep, st["n_det"])
    np.testing.assert_array_equal(ch.n_gt.cpu().numpy(), st["n_gt"])
    ks = ch.rec_k.cpu().numpy()
    fl = ch.rec_flags.cpu().numpy().view(np.uint64) & np.uint64((1 << 56) - 1)
    for i in range(n):
        mine = {int(ks[i, q]): int(fl[i, q]) for q in range(n_keep[i])}
        want = {int(st["det_k"][i, q]): int(ref["flags"][i, q]) for q in range(st["n_det"][i])}
        assert mine == want, f"image {i}"
    got = m.compute()
    want = results.assemble(ref["ap"], ref["overall"], ref["sizes"], list(thr), 20, n)
    assert sorted(got) == sorted(want)
    for k in want:
        assert float(got[k]) == pytest.approx(float(want[k]), rel=1e-9, abs=1e-12), k
    return m


def _dense_scene(n, seed, n_cls):
    """Every cell holds a ground truth; every candidate is confident; only n_cls classes occur."""
    g = torch.Generator().manual_seed(seed)
    tgt = torch.zeros(n, 7, 7, 30)
    tgt[..., 0:2] = torch.rand(n, 7, 7, 2, generator=g)
    tgt[..., 2:4] = torch.rand(n, 7, 7, 2, generator=g) * 0.3 + 0.02
    tgt[..., 4] = 1.0
    cls = torch.randint(0, n_cls, (n, 7, 7), generator=g)
    tgt[..., 10:].scatter_(-1, cls[..., None], 1.0)
    pred = torch.rand(n, 7, 7, 30, generator=g) * 0.2
    pred[..., 0:4] = tgt[..., 0:4] + torch.randn(n, 7, 7, 4, generator=g) * 0.02
    pred[..., 5:9] = tgt[..., 0:4] + torch.randn(n, 7, 7, 4, generator=g) * 0.05
    pred[..., 4] = torch.rand(n, 7, 7, generator=g) * 0.5 + 0.5
    pred[..., 9] = torch.rand(n, 7, 7, generator=g) * 0.5 + 0.5
    pcls = torch.where(torch.rand(n, 7, 7, generator=g) < 0.8, cls, torch.randint(0, n_cls, (n, 7, 7), generator=g))
    pred[..., 10:].scatter_(-1, pcls[..., None], 0.9)
    return pred, tgt


@pytest.mark.parametrize("n_cls", [1, 2, 20])
def test_dense_scenes_49_ground_truths(n_cls):
    pred, tgt = _dense_scene(48, 10 + n_cls, n_cls)
    _check(pred, tgt, 0.05)
    _check(pred, tgt, 0.05, nms=0.95)   # almost nothing suppressed: up to 98 kept of one class


def test_duplicate_boxes_share_a_ground_truth():
    # many identical detections claim the same ground truth: only the first (stable order) may be a TP
    pred, tgt = _dense_scene(16, 3, 1)
    pred[..., 5:9] = pred[..., 0:4]
    pred[..., 9] = pred[..., 4]
    _check(pred, tgt, 0.05, nms=1.5)    # threshold above any IoU: NMS keeps all 98


@pytest.mark.parametrize("T", [1, 2, 7, 13])
def test_threshold_counts(T):
    thr = [0.3 + 0.05 * i for i in range(T)]
    tgt = synth.make_targets(200, seed=5, p_obj=0.2, size_mix=True)
    pred = synth.make_preds(200, "correlated", seed=5, targets=tgt)
    _check(pred, tgt, 0.01, thr=thr)


def test_unsorted_and_repeated_thresholds():
    tgt = synth.make_targets(128, seed=6, p_obj=0.2, size_mix=True)
    pred = synth.make_preds(128, "correlated", seed=6, targets=tgt)
    _check(pred, tgt, 0.01, thr=[0.9, 0.5, 0.5, 0.75, 0.1])


def test_nan_and_inf_values():
    tgt = synth.make_targets(64, seed=8, p_obj=0.3, size_mix=True)
    pred = synth.make_preds(64, "correlated", seed=8, targets=tgt)
    g = torch.Generator().manual_seed(8)
    hit = torch.rand(pred.shape, generator=g)
    pred[hit < 0.01] = float("nan")
    pred[(hit >= 0.01) & (hit < 0.015)] = float("inf")
    pred[(hit >= 0.015) & (hit < 0.02)] = float("-inf")
    _check(pred, tgt, 0.01)


def test_rawhead_negative_sizes_and_zero_area():
    tgt = synth.make_targets(256, seed=9, p_obj=0.2, size_mix=True)
    pred = synth.make_preds(256, "rawhead", seed=9, targets=tgt)
    pred[:64, :, :, 2] = 0.0            # zero-width boxes: union can be 0 (metrics.py:337-338)
    tgt[:32, :, :, 2] *= 0.0
    _check(pred, tgt, 0.01)


def test_unaligned_views_are_accepted():
    tgt = synth.make_targets(33, seed=4, p_obj=0.2, size_mix=True)
    pred = synth.make_preds(33, "correlated", seed=4, targets=tgt)
    buf = torch.zeros(pred.numel() + 1, device=DEV)
    buf[1:] = pred.to(DEV).reshape(-1)
    view = buf[1:].view_as(pred)        # 4-byte aligned only
    a = mAPMetric(num_classes=20)
    a.update(view, tgt.to(DEV))
    b = mAPMetric(num_classes=20)
    b.update(pred.to(DEV), tgt.to(DEV))
    assert a.compute() == b.compute()


def test_full_size_invariants():
    n = 65536
    tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(n, "uniform", seed=1234, targets=tgt)
    gt_count = torch.zeros(80, dtype=torch.int32, device=DEV)
    ch = head.map_update(pred.to(DEV), tgt.to(DEV), 0.3, 0.4, DEFAULT_THR, gt_count)
    dec = head.decode_nms(pred.to(DEV), 0.3, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, want_boxes=False)
    assert torch.equal(ch.n_keep, dec.n_keep)                       # fused update == stand-alone decode+NMS
    assert bool((dec.n_keep <= dec.n_cand).all())                   # kept is a subset of the candidates
    n_keep = ch.n_keep.cpu().numpy()
    valid = np.arange(98)[None, :] < n_keep[:, None]
    ks_f = np.sort(np.where(valid, ch.rec_k.cpu().numpy(), 255), axis=1)
    ks_d = np.sort(np.where(valid, dec.keep_k.cpu().numpy(), 255), axis=1)
    assert np.array_equal(ks_f, ks_d)                               # same kept set per image
    fl = ch.rec_flags.cpu().numpy().view(np.uint64)
    cls = (fl >> np.uint64(56)).astype(np.int64)
    gtc = gt_count.cpu().numpy().reshape(4, 20)
    assert gtc[0].sum() == int(ch.n_gt.sum()) and np.array_equal(gtc[0], gtc[1:].sum(0))
    for s in range(4):                                              # TPs never exceed the ground truths
        for t in (0, 9):
            bit = ((fl >> np.uint64(s * 10 + t)) & np.uint64(1)).astype(bool) & valid
            tp_per_class = np.bincount(cls[bit], minlength=20)
            assert (tp_per_class <= gtc[s]).all(), (s, t)
    # a TP at a higher threshold is a TP at every lower one only per ground truth, but the count is monotone
    tp_lo = int((((fl >> np.uint64(0)) & np.uint64(1)).astype(bool) & valid).sum())
    tp_hi = int((((fl >> np.uint64(9)) & np.uint64(1)).astype(bool) & valid).sum())
    assert tp_hi <= tp_lo
    # sorted keys inside each image's class groups: (class asc, confidence desc)
    key = ch.rec_key.cpu().numpy().view(np.uint64)
    same = valid[:, 1:] & (cls[:, 1:] == cls[:, :-1])
    assert (key[:, 1:][same] >= key[:, :-1][same]).all()
    assert (cls[:, 1:][valid[:, 1:]] >= cls[:, :-1][valid[:, 1:]]).all()


def _by_candidate(n_keep, ks, vals, fill):
    """(N,98) array indexed by candidate k from per-position arrays (positions >= n_keep ignored)."""
    n = len(n_keep)
    live = np.arange(98)[None, :] < n_keep[:, None]
    out = np.full((n, 98), fill, dtype=vals.dtype)
    rows = np.repeat(np.arange(n), 98).reshape(n, 98)
    out[rows[live], ks.astype(np.int64)[live]] = vals[live]
    return out


def test_full_size_oracle_parity_config1():
    """BASELINE.json configs[1] at its stated size: 65,536 `uniform` images, conf 0.3 / NMS IoU 0.4.
    Ordered keep lists of both flavours ((metrics IoU, class-grouped) and (inference IoU, global
    confidence order)) against the C oracle, and the fused update's TP flag words against the oracle's
    compute(), every image, bit for bit."""
    import os

    n = 65536
    th = min(os.cpu_count() or 1, 32)
    tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(n, "uniform", seed=1234, targets=tgt)
    pd, td = pred.to(DEV), tgt.to(DEV)
    live_of = lambda nk: np.arange(98)[None, :] < nk[:, None]  # noqa: E731
    for iou_mode, order, o_iou, o_order in ((head.IOU_METRICS, head.ORDER_CLASS_GROUPED, c_oracle.IOU_METRICS, c_oracle.ORDER_CLASS_GROUPED),
                                            (head.IOU_INFER, head.ORDER_GLOBAL_CONF, c_oracle.IOU_INFER, c_oracle.ORDER_GLOBAL_CONF)):
        st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), 0.3, 0.4, o_iou, o_order, n_threads=th)
        r = head.decode_nms(pd, 0.3, 0.4, iou_mode, order, want_boxes=False)
        nk = r.n_keep.cpu().numpy()
        np.testing.assert_array_equal(nk, st["n_det"])
        live = live_of(nk)
        assert np.array_equal(np.where(live, r.keep_k.cpu().numpy().astype(np.int32), -1), np.where(live, st["det_k"], -1))
        assert np.array_equal(np.where(live, r.keep_cls.cpu().numpy().astype(np.int32), -1), np.where(live, st["det_cls"], -1))
        assert np.where(live, r.keep_conf.cpu().numpy(), 0.0).tobytes() == np.where(live, st["det_conf"], 0.0).tobytes()
        if order == head.ORDER_CLASS_GROUPED:
            ref = c_oracle.compute(st, DEFAULT_THR, n_threads=th)
            gt_count = torch.zeros(80, dtype=torch.int32, device=DEV)
            ch = head.map_update(pd, td, 0.3, 0.4, DEFAULT_THR, gt_count)
            np.testing.assert_array_equal(ch.n_keep.cpu().numpy(), st["n_det"])
            np.testing.assert_array_equal(ch.n_gt.cpu().numpy(), st["n_gt"])
            fl = ch.rec_flags.cpu().numpy().view(np.uint64) & np.uint64((1 << 56) - 1)
            mine = _by_candidate(nk, ch.rec_k.cpu().numpy(), fl, np.uint64(0))
            want = _by_candidate(nk, st["det_k"], ref["flags"].astype(np.uint64), np.uint64(0))
            assert np.array_equal(mine, want)
            assert int((mine & np.uint64(1)).sum()) > 0            # some true positives exist at IoU 0.5


# ----------------------------------------------------------------------------- sort plan / AP reduce on raw records
def _ap_from_records(key, flags, gt_count, C, T):
    """numpy restatement of metrics.py:382-436 over (key, flags) records: stable sort per class by key
    (ascending key = descending confidence), cumulative TP, precision / recall, 11-point AP."""
    ap = np.zeros((4, C, T))
    cls = (flags >> np.uint64(56)).astype(np.int64)
    levels = np.linspace(0, 1, 11)
    for c in range(C):
        idx = np.nonzero(cls == c)[0]
        order = idx[np.argsort(key[idx], kind="stable")]
        f = flags[order]
        for s in range(4):
            ngt = int(gt_count[s * C + c])
            if ngt == 0 or len(order) == 0:
                continue
            for t in range(T):
                tp = ((f >> np.uint64(s * T + t)) & np.uint64(1)).astype(np.float64)
                tpc, fpc = np.cumsum(tp), np.cumsum(1 - tp)
                prec = np.concatenate([[1.0], tpc / (tpc + fpc + 1e-6)])
                rec = np.concatenate([[0.0], tpc / ngt])
                a = 0.0
                for lv in levels:
                    a += (prec[rec >= lv].max() if (rec >= lv).any() else 0.0) / 11
                ap[s, c, t] = a
    return ap


def _sort_ctl(dev, m, C, T):
    """(overflow, queued runs) of the last yh_ap_reduce on this device: SortCtl sits at the head of the
    workspace (csrc/ap_reduce.cu: tile_counter[16], err, overflow, key_or, key_nand, work_count)."""
    ws = head._AP_WS.get(torch.device(dev), _lib.lib().yh_ap_workspace_bytes(int(m), C, T))
    w = ws[:96].cpu().numpy().view(np.uint32)
    return int(w[17]), int(w[22])


# short plan (four passes over the composite's leading 32 bits + fix_runs) and its fallback:
#   short_sparse    50 varying key bits, no two records agree in the leading bits
#   short_runs      128 leading values per class, random tails: every run is out of order, all shorter than 32
#   short_equal     long runs of IDENTICAL keys next to ordinary ones: nothing to fix, no fallback
#   short_clusters  4 leading values per class, random tails: runs of thousands -> radix_fallback
#   short_mixed     mostly sparse, one long unordered run -> radix_fallback
SHORT_KINDS = ["short_sparse", "short_runs", "short_equal", "short_clusters", "short_clusters42", "short_mixed"]


def _check_sorted_arrays(key, flags, k_dev, f_dev, fell_back):
    """The sort ping-pongs between the caller's arrays and the workspace; when the number of passes it ran is
    even, the caller's arrays hold the result and must equal numpy's stable sort by (class, key) record for
    record (the AP table alone would forgive a few misplaced neighbours)."""
    vary = int(np.bitwise_or.reduce(key) & np.bitwise_or.reduce(~key))
    nb = vary.bit_length() - ((vary & -vary).bit_length() - 1) if vary else 0
    full = (nb + 5 + 7) // 8
    passes = full if (nb > 59 or nb + 5 <= 32 or fell_back) else 4
    if passes % 2:
        return
    cls = (flags >> np.uint64(56)).astype(np.int64)
    order = np.lexsort((np.arange(len(key)), key, cls))
    got_k, got_f = k_dev.cpu().numpy().view(np.uint64), f_dev.cpu().numpy().view(np.uint64)
    bad = np.nonzero((got_k != key[order]) | (got_f != flags[order]))[0]
    assert len(bad) == 0, (len(bad), bad[:8], [hex(int(x)) for x in got_k[bad[:4]]], [hex(int(x)) for x in key[order][bad[:4]]])


# the same at a size where radix_fallback's persistent CTAs each take several tiles per pass and its grid
# barrier spans every SM (1.5 M records = 733 tiles): the two-cluster keys (6 passes, result in the caller's
# arrays) and full 64-bit keys (9 passes, everything left to the fallback); and the short plan with ~2,000 runs
BIG_KINDS = ["big:short_clusters42", "big:full64", "big:short_sparse"]


@pytest.mark.parametrize("kind", ["full64", "equal", "narrow", "high_bits", "two_values", "one_record", "tile_edge"]
                         + SHORT_KINDS + BIG_KINDS)
def test_ap_reduce_sort_plans(kind):
    rng = np.random.default_rng(len(kind) * 131 + ord(kind[0]))
    C, T = 20, 3
    m = {"one_record": 1, "tile_edge": 2048 * 3, "full64": 70_001}.get(kind, 30_000)
    if kind.startswith("big:"):
        kind, m = kind[4:], 1_500_000
    if kind == "full64":
        key = rng.integers(0, 2**64, size=m, dtype=np.uint64)
    elif kind == "equal":
        key = np.full(m, 0x3FE0_0000_0000_0000, np.uint64)
    elif kind == "narrow":
        key = np.uint64(0x4000_0000_0000_0000) + (rng.integers(0, 1 << 13, size=m).astype(np.uint64) << np.uint64(20))
    elif kind == "high_bits":
        key = rng.integers(0, 4, size=m).astype(np.uint64) << np.uint64(62)
    elif kind == "two_values":
        key = np.where(rng.random(m) < 0.5, np.uint64(5), np.uint64(1 << 40)).astype(np.uint64)
    elif kind == "short_sparse":
        key = np.uint64(0x3FD0_0000_0000_0000) + rng.integers(0, 1 << 50, size=m).astype(np.uint64)
    elif kind == "short_runs":
        key = (rng.integers(0, 128, size=m).astype(np.uint64) << np.uint64(42)) | rng.integers(0, 1 << 20, size=m).astype(np.uint64)
    elif kind == "short_equal":
        key = np.uint64(0x3FD0_0000_0000_0000) + rng.integers(0, 1 << 50, size=m).astype(np.uint64)
        key[rng.random(m) < 0.5] = np.uint64(0x3FD0_0000_0000_0000) + np.uint64(12345 << 20)
    elif kind == "short_clusters":      # 50 varying bits: the fallback runs 7 passes
        key = (rng.integers(0, 4, size=m).astype(np.uint64) << np.uint64(48)) | rng.integers(0, 1 << 20, size=m).astype(np.uint64)
    elif kind == "short_clusters42":    # 42 varying bits: 6 passes, the result lands in the caller's arrays
        key = (rng.integers(0, 4, size=m).astype(np.uint64) << np.uint64(40)) | rng.integers(0, 1 << 15, size=m).astype(np.uint64)
    elif kind == "short_mixed":
        key = np.uint64(0x3FD0_0000_0000_0000) + rng.integers(0, 1 << 42, size=m).astype(np.uint64)
        key[:200] = np.uint64(0x3FD0_0000_0000_0000) + np.uint64(777 << 15) + rng.integers(0, 1 << 15, size=200).astype(np.uint64)
    else:
        key = rng.integers(0, 2**64, size=m, dtype=np.uint64)
    cls = rng.integers(0, C, size=m).astype(np.uint64)
    if kind == "short_mixed":
        cls[:200] = 7
    bits = rng.integers(0, 1 << (4 * T + 1), size=m).astype(np.uint64)
    flags = (cls << np.uint64(56)) | bits
    gt_count = rng.integers(0, 200, size=4 * C).astype(np.int32) + 400
    gt_count[3] = 0
    want = _ap_from_records(key, flags, gt_count, C, T)
    k = torch.from_numpy(key.view(np.int64).copy()).to(DEV)
    f = torch.from_numpy(flags.view(np.int64).copy()).to(DEV)
    g = torch.from_numpy(gt_count).to(DEV)
    ap, totals, packed = head.ap_reduce(k, f, m, g, C, T)
    np.testing.assert_allclose(ap.cpu().numpy(), want, rtol=1e-12, atol=0)
    assert int(totals[1]) == m and int(totals[0]) == int(((bits >> np.uint64(4 * T)) & np.uint64(1)).sum())
    overflow, queued = _sort_ctl(g.device, m, C, T)
    _check_sorted_arrays(key, flags, k, f, overflow != 0)
    if kind in SHORT_KINDS:  # the path each case is meant to take
        assert overflow == (1 if kind in ("short_clusters", "short_clusters42", "short_mixed") else 0), (overflow, queued)
        if kind == "short_runs":
            assert queued > 1000
        if kind in ("short_sparse", "short_equal") and m <= 30_000:
            assert queued == 0
    # same records, count supplied from the device with a larger capacity
    cap = m + 5000
    k2 = torch.zeros(cap, dtype=torch.int64, device=DEV); k2[:m] = torch.from_numpy(key.view(np.int64).copy()).to(DEV)
    f2 = torch.zeros(cap, dtype=torch.int64, device=DEV); f2[:m] = torch.from_numpy(flags.view(np.int64).copy()).to(DEV)
    ap2, _, packed2 = head.ap_reduce(k2, f2, cap, g, C, T, m_dev=torch.tensor([m], dtype=torch.int64, device=DEV))
    assert torch.equal(ap2, ap)
    a, tp, n, gc = head.unpack_ap(packed2, C, T)
    assert n == m and np.array_equal(gc.reshape(-1), gt_count) and np.array_equal(a, ap.cpu().numpy())


# ----------------------------------------------------------------------------- round-2 regressions
def test_second_backward_over_a_retained_graph_recomputes_the_gradient():
    """The reference's autograd graph can be walked twice with retain_graph=True; the fused loss gives its
    gradient buffer away on the first backward and re-runs the kernel for any further one."""
    from oracle import loss_oracle
    from yolo_v1_b200 import YOLOLoss

    tgt = synth.make_targets(8, seed=3, p_obj=0.2)
    pred = synth.make_preds(8, "uniform", seed=4)
    p = pred.to(DEV).requires_grad_(True)
    total, _ = YOLOLoss()(p, tgt.to(DEV))
    total.backward(retain_graph=True)
    g1 = p.grad.clone()
    p.grad = None
    (total * 3.0).backward()
    _, ref = loss_oracle.yolo_loss_and_grad(pred, tgt)
    np.testing.assert_allclose(g1.cpu().numpy(), ref.numpy(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(p.grad.cpu().numpy(), 3.0 * ref.numpy(), rtol=1e-5, atol=1e-7)


def test_encode_target_without_a_device_argument():
    from oracle import encode_oracle
    from yolo_v1_b200 import targets

    boxes, cls = [[0.3, 0.4, 0.2, 0.1], [0.8, 0.1, 0.05, 0.3]], [3, 17]
    got = targets.encode_target(boxes, cls)
    assert got.is_cuda and got.cpu().numpy().tobytes() == encode_oracle.encode_target(boxes, cls).tobytes()
    assert targets.encode_targets([boxes, []], [cls, []]).shape == (2, 7, 7, 30)


def test_concurrent_streams_do_not_share_scratch():
    """Two streams run loss + update + compute at the same time on different data (stream-keyed workspaces,
    ticket ring): each must give the result it gives alone."""
    from yolo_v1_b200 import YOLOLoss

    sets = []
    for seed in (11, 12):
        tgt = synth.make_targets(6000, seed=seed, p_obj=0.055, size_mix=True)
        pred = synth.make_preds(6000, "correlated", seed=seed, targets=tgt)
        sets.append((pred.to(DEV), tgt.to(DEV)))

    def run(p, t):
        total, parts = YOLOLoss(defer_loss_dict=True)(p.clone().requires_grad_(True), t)
        m = mAPMetric(num_classes=20, conf_threshold=0.05)
        m.update(p, t)
        packed = m._enqueue_reduce()
        return parts._parts, packed

    alone = [tuple(x.clone() for x in run(p, t)) for p, t in sets]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for _ in range(5):
        outs = []
        for st, (p, t) in zip(streams, sets):
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                outs.append(run(p, t))
        torch.cuda.synchronize()
        for (a_parts, a_packed), (b_parts, b_packed) in zip(alone, outs):
            assert torch.equal(a_parts, b_parts) and torch.equal(a_packed, b_packed)


def test_one_process_every_visible_device():
    """The ticket array is a __device__ symbol with one address per device, and kernel attributes are per
    device: a process that uses cuda:0, then cuda:1, then cuda:0 again must get the same answer everywhere
    (also the static-scratch loss, whose array address is looked up per device).  On a one-GPU box the three
    passes all run on cuda:0 -- still a repeatability check; the two-device form ran on a 2-GPU box in round 2."""
    from yolo_v1_b200 import _lib

    last = torch.cuda.device_count() - 1
    devices = ["cuda:0", f"cuda:{min(1, last)}", "cuda:0"]
    tgt = synth.make_targets(3000, seed=21, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(3000, "correlated", seed=21, targets=tgt)
    res = []
    for d in devices:
        m = mAPMetric(num_classes=20, conf_threshold=0.05)
        m.update(pred.to(d), tgt.to(d))
        res.append(m.compute())
        r = head.decode_nms(pred.to(d), 0.3, 0.4, head.IOU_INFER, head.ORDER_GLOBAL_CONF, want_boxes=False)
        live = torch.arange(98, device=d)[None, :] < r.n_keep[:, None]      # entries past n_keep are uninitialised
        res.append(torch.where(live, r.keep_k, torch.zeros_like(r.keep_k)).cpu())
        p_d, t_d = pred.to(d), tgt.to(d)
        parts = torch.empty(5, dtype=torch.float32, device=d)
        with torch.cuda.device(d):
            rc = _lib.lib().yh_loss_fwd_bwd_static(p_d.data_ptr(), 0, t_d.data_ptr(), 3000, 7, 2, 20, 5.0, 0.5,
                                                   parts.data_ptr(), None, _lib.stream_ptr(torch.device(d)))
        assert rc == 0
        res.append(parts.cpu())
    assert res[0] == res[3] == res[6]
    assert torch.equal(res[1], res[4]) and torch.equal(res[1], res[7])
    assert torch.equal(res[2], res[5]) and torch.equal(res[2], res[8])
