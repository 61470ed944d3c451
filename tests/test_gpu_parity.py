"""GPU parity: the CUDA kernels (through the C ABI / Python host) against
  (a) golden vectors produced by the unmodified reference (tests/golden/*.npz), and
  (b) the CPU oracle (oracle/) on the same seeded inputs at larger sizes.
Bars (BASELINE.json north_star): NMS keep indices and TP/FP flags bit-exact; loss,
gradients and AP within 1e-5 relative (fp32)."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, loss_oracle, results
from yolo_v1_b200 import YOLOLoss, head, mAPMetric, synth

pytestmark = pytest.mark.gpu

DEV = "cuda"
LOSS_CASES = ["uniform1", "uniform2", "uniform16", "uniform64", "noobj2", "bothslots8", "slot1only4",
              "ioutie4", "rawhead8", "ties8", "correlated16", "clamp4"]
KEYS = ("total", "coord", "conf_obj", "conf_noobj", "class")


def _grad_close(got, ref, rtol=1e-5):
    scale = max(1.0, float(np.abs(ref).max()))
    np.testing.assert_allclose(got, ref, rtol=rtol, atol=1e-7 * scale)


# ------------------------------------------------------------------------- loss
@pytest.mark.parametrize("case", LOSS_CASES)
def test_loss_matches_reference_golden(golden, case):
    g = golden("loss")
    pred = torch.from_numpy(g[f"{case}/pred"]).to(DEV).requires_grad_(True)
    tgt = torch.from_numpy(g[f"{case}/tgt"]).to(DEV)
    total, d = YOLOLoss()(pred, tgt)
    total.backward()
    got = np.array([d[k] for k in KEYS])
    np.testing.assert_allclose(got, g[f"{case}/parts"], rtol=1e-5, atol=1e-9)
    assert float(total) == pytest.approx(d["total"], rel=1e-7)
    _grad_close(pred.grad.cpu().numpy(), g[f"{case}/grad"])


def test_loss_upstream_gradient_scaling(golden):
    g = golden("loss")
    pred = torch.from_numpy(g["uniform16/pred"]).to(DEV).requires_grad_(True)
    tgt = torch.from_numpy(g["uniform16/tgt"]).to(DEV)
    total, _ = YOLOLoss()(pred, tgt)
    (total * 1024.0).backward()
    _grad_close(pred.grad.cpu().numpy(), g["uniform16_scaled/grad"])


@pytest.mark.parametrize("n,seed", [(256, 1), (1000, 2), (4952, 3)])
def test_loss_matches_oracle_seeded(n, seed):
    tgt = synth.make_targets(n, seed=seed, p_obj=0.06, both_slots=0.1)
    pred = synth.make_preds(n, "uniform", seed=seed + 50)
    parts_ref, grad_ref = loss_oracle.yolo_loss_and_grad(pred, tgt)
    p = pred.to(DEV).requires_grad_(True)
    total, d = YOLOLoss()(p, tgt.to(DEV))
    total.backward()
    np.testing.assert_allclose(np.array([d[k] for k in KEYS]), parts_ref.double().numpy(), rtol=1e-5)
    _grad_close(p.grad.cpu().numpy(), grad_ref.numpy())


def test_loss_no_grad_and_determinism():
    tgt = synth.make_targets(300, seed=5).to(DEV)
    pred = synth.make_preds(300, "uniform", seed=6).to(DEV)
    crit = YOLOLoss()
    with torch.no_grad():
        t1, d1 = crit(pred, tgt)
    assert not t1.requires_grad
    p = pred.clone().requires_grad_(True)
    t2, d2 = crit(p, tgt)
    t2.backward()
    g1 = p.grad.clone()
    p.grad = None
    t3, d3 = crit(p, tgt)
    t3.backward()
    assert d1 == d2 == d3                      # fixed reduction order: bitwise repeatable
    assert torch.equal(g1, p.grad)


@pytest.mark.parametrize("dt,name", [(torch.float16, "f16"), (torch.bfloat16, "bf16")])
def test_loss_half_precision_predictions(golden, dt, name):
    # The kernel widens half predictions to fp32; the reference does part of its arithmetic
    # in the narrow type, so this is a looser, documented tolerance.
    g = golden("loss")
    pred = torch.from_numpy(g[f"uniform16_{name}/pred_as_f32"]).to(DEV).to(dt).requires_grad_(True)
    tgt = torch.from_numpy(g["uniform16/tgt"]).to(DEV)
    total, d = YOLOLoss()(pred, tgt)
    total.backward()
    assert pred.grad.dtype == dt
    tol = 2e-3 if dt == torch.float16 else 2e-2
    np.testing.assert_allclose(np.array([d[k] for k in KEYS]), g[f"uniform16_{name}/parts"], rtol=tol)
    ref = g[f"uniform16_{name}/grad_as_f32"]
    # compare where the responsible-box choice cannot have flipped: overall relative error
    err = np.abs(pred.grad.float().cpu().numpy() - ref).sum() / np.abs(ref).sum()
    assert err < 5 * tol


@pytest.mark.parametrize("case", LOSS_CASES)
def test_update_with_loss_matches_reference_golden_and_the_two_separate_kernels(golden, case):
    """yh_map_update_loss: the loss rides on the update kernel.  Loss parts vs the reference's (rtol 1e-5), gradient
    bit-identical to the stand-alone loss kernel's (same per-cell arithmetic), records identical to yh_map_update's."""
    g = golden("loss")
    pred, tgt = torch.from_numpy(g[f"{case}/pred"]).to(DEV), torch.from_numpy(g[f"{case}/tgt"]).to(DEV)
    m = mAPMetric(num_classes=20, conf_threshold=0.05)
    parts, grad = m.update_with_loss(pred, tgt, YOLOLoss())
    np.testing.assert_allclose(parts.cpu().numpy().astype(np.float64), g[f"{case}/parts"], rtol=1e-5, atol=1e-9)
    ref = g[f"{case}/grad"]
    np.testing.assert_allclose(grad.cpu().numpy(), ref, rtol=1e-5, atol=1e-7 * max(1.0, np.abs(ref).max()))
    p2 = pred.clone().requires_grad_(True)
    total, _ = YOLOLoss()(p2, tgt)
    total.backward()
    assert torch.equal(grad, p2.grad)
    m2 = mAPMetric(num_classes=20, conf_threshold=0.05)
    m2.update(pred, tgt)
    a, b = m._chunks[0], m2._chunks[0]
    assert torch.equal(a.n_keep, b.n_keep) and torch.equal(a.n_gt, b.n_gt) and torch.equal(m._gt_count, m2._gt_count)
    live = (torch.arange(98, device=DEV)[None, :] < a.n_keep[:, None])
    for x, y in ((a.rec_key, b.rec_key), (a.rec_flags, b.rec_flags), (a.rec_k, b.rec_k)):
        assert torch.equal(x[live], y[live])
    assert m.compute() == m2.compute()
    fwd_parts, none = mAPMetric(num_classes=20).update_with_loss(pred, tgt, YOLOLoss(), backward=False)
    assert none is None and torch.equal(fwd_parts, parts)          # forward only: same sums, repeatable


def test_loss_rejects_cpu_and_bad_shapes():
    crit = YOLOLoss()
    with pytest.raises(RuntimeError):
        crit(torch.zeros(1, 7, 7, 30), torch.zeros(1, 7, 7, 30))
    with pytest.raises(ValueError):
        crit(torch.zeros(1, 7, 7, 29, device=DEV), torch.zeros(1, 7, 7, 29, device=DEV))
    # other head shapes run the generic kernels (tests/test_generic_shapes.py): S = 14 is fine for the loss, which has
    # no candidate limit; more than 8 boxes per cell is outside the build
    total, parts = YOLOLoss(S=14)(torch.zeros(1, 14, 14, 30, device=DEV), torch.zeros(1, 14, 14, 30, device=DEV))
    assert float(total) == 0.0 and parts["conf_noobj"] == 0.0
    with pytest.raises(NotImplementedError):
        YOLOLoss(B=9)(torch.zeros(1, 7, 7, 65, device=DEV), torch.zeros(1, 7, 7, 65, device=DEV))


# ------------------------------------------------------------------ decode + NMS
PROFILES = ["uniform", "realistic", "ties", "rawhead", "correlated"]


def _assert_keep_equal(r, n_keep, keep_k, what):
    got_n = r.n_keep.cpu().numpy()
    np.testing.assert_array_equal(got_n, n_keep, err_msg=what)
    gk = r.keep_k.cpu().numpy()
    for i in range(len(n_keep)):
        np.testing.assert_array_equal(gk[i, :n_keep[i]], keep_k[i, :n_keep[i]], err_msg=f"{what} image {i}")


@pytest.mark.parametrize("profile", PROFILES)
@pytest.mark.parametrize("conf_thr", [0.01, 0.3, 0.5])
def test_decode_nms_matches_reference_golden(golden, profile, conf_thr):
    g = golden("decode_nms")
    pred = torch.from_numpy(g[f"{profile}/pred"]).to(DEV)
    tag = f"{profile}/c{conf_thr}"
    r = head.decode_nms(pred, conf_thr, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED)
    gk = g[f"{tag}/metrics_keep_k"]
    n_keep = (gk >= 0).sum(1)
    _assert_keep_equal(r, n_keep, gk, "metrics order")
    conf, cls, box = r.keep_conf.cpu().numpy(), r.keep_cls.cpu().numpy(), r.keep_box.cpu().numpy()
    for i in range(pred.shape[0]):
        n = n_keep[i]
        assert conf[i, :n].tobytes() == g[f"{tag}/metrics_conf"][i, :n].tobytes()
        np.testing.assert_array_equal(cls[i, :n], g[f"{tag}/metrics_cls"][i, :n])
        assert box[i, :n].tobytes() == g[f"{tag}/metrics_box"][i, :n].tobytes()
    r2 = head.decode_nms(pred, conf_thr, 0.4, head.IOU_INFER, head.ORDER_GLOBAL_CONF)
    ik, raises = g[f"{tag}/infer_keep_k"], g[f"{tag}/infer_raises"]
    bad = r2.bad_range.cpu().numpy().astype(bool)
    np.testing.assert_array_equal(bad, raises)   # where the reference's pydantic schema raised
    ok = ~raises
    n2 = (ik >= 0).sum(1)
    got_n, got_k = r2.n_keep.cpu().numpy(), r2.keep_k.cpu().numpy()
    for i in np.nonzero(ok)[0]:
        assert got_n[i] == n2[i]
        np.testing.assert_array_equal(got_k[i, :n2[i]], ik[i, :n2[i]])


@pytest.mark.parametrize("profile", PROFILES)
@pytest.mark.parametrize("mode,order", [(0, 0), (1, 1), (0, 1), (1, 0)])
def test_decode_nms_matches_oracle_seeded(profile, mode, order):
    n = 1500
    tgt = synth.make_targets(n, seed=77, p_obj=0.08, size_mix=True)
    pred = synth.make_preds(n, profile, seed=78, targets=tgt)
    conf_thr = 0.3 if profile in ("uniform", "ties") else 0.05
    st = c_oracle.update_batch(pred.numpy(), None, conf_thr, 0.4, mode, order, n_threads=8)
    r = head.decode_nms(pred.to(DEV), conf_thr, 0.4, mode, order)
    _assert_keep_equal(r, st["n_det"], st["det_k"], f"{profile} mode {mode} order {order}")
    conf = r.keep_conf.cpu().numpy()
    box = r.keep_box.cpu().numpy()
    for i in range(0, n, 97):
        k = st["n_det"][i]
        assert conf[i, :k].tobytes() == st["det_conf"][i, :k].tobytes()
        assert box[i, :k].tobytes() == st["det_box"][i, :k].tobytes()


def test_decode_nms_edge_cases():
    # nothing above the threshold; everything one class and identical boxes; single image (3-D input)
    z = torch.zeros(3, 7, 7, 30, device=DEV)
    r = head.decode_nms(z, 0.01, 0.4)
    assert r.n_keep.tolist() == [0, 0, 0]
    one = torch.zeros(7, 7, 30)
    one[..., 0:4] = torch.tensor([0.5, 0.5, 0.2, 0.2])
    one[..., 5:9] = torch.tensor([0.5, 0.5, 0.2, 0.2])
    one[..., 4] = 0.9
    one[..., 9] = 0.8
    one[..., 10] = 1.0
    st = c_oracle.update_batch(one[None].numpy(), None, 0.1, 0.4)
    r = head.decode_nms(one.to(DEV), 0.1, 0.4)
    _assert_keep_equal(r, st["n_det"], st["det_k"], "98 same-class candidates")
    # all 98 candidates of one class, pairwise disjoint -> 98 kept, 98 NMS rounds
    far = torch.zeros(1, 7, 7, 30)
    far[..., 0:2] = 0.25
    far[..., 5:7] = 0.75
    far[..., 2:4] = 0.01
    far[..., 7:9] = 0.01
    far[..., 4] = torch.linspace(0.5, 0.9, 49).reshape(7, 7)
    far[..., 9] = torch.linspace(0.4, 0.45, 49).reshape(7, 7)
    far[..., 12] = 1.0
    st = c_oracle.update_batch(far.numpy(), None, 0.1, 0.4)
    r = head.decode_nms(far.to(DEV), 0.1, 0.4)
    assert int(r.n_keep[0]) == 98
    _assert_keep_equal(r, st["n_det"], st["det_k"], "98 kept")


# ----------------------------------------------------------------------- mAP
MAP_CASES = ["corr64", "ties48", "uniform32", "rawhead24", "corr40_thr3", "dupties48"]


def _flags_by_candidate(n_keep, ks, flags):
    out = []
    for i in range(len(n_keep)):
        out.append({int(ks[i, q]): int(flags[i, q]) for q in range(n_keep[i])})
    return out


def _fused_flags(metric):
    ch = metric._chunks[0]
    n_keep = ch.n_keep.cpu().numpy()
    fl = ch.rec_flags.cpu().numpy().view(np.uint64) & np.uint64((1 << 56) - 1)
    return n_keep, _flags_by_candidate(n_keep, ch.rec_k.cpu().numpy(), fl)


@pytest.mark.parametrize("case", MAP_CASES)
def test_map_matches_reference_golden(golden, case):
    g = golden("map_small")
    pred, tgt = torch.from_numpy(g[f"{case}/pred"]), torch.from_numpy(g[f"{case}/tgt"])
    thr = [float(t) for t in g[f"{case}/thresholds"]]
    T = len(thr)
    metric = mAPMetric(num_classes=20, iou_thresholds=thr, conf_threshold=float(g[f"{case}/conf_thr"]))
    metric.update(pred.to(DEV), tgt.to(DEV))
    # TP/FP flags, bit-exact, keyed by candidate index
    gk, gflags = g[f"{case}/keep_k"], g[f"{case}/flags"]
    n_ref = (gk >= 0).sum(1)
    n_keep, mine = _fused_flags(metric)
    np.testing.assert_array_equal(n_keep, n_ref)
    ap_bits = (1 << (4 * T)) - 1
    for i in range(pred.shape[0]):
        ref = {int(gk[i, q]): int(gflags[i, q]) for q in range(n_ref[i])}
        assert {k: v & ap_bits for k, v in mine[i].items()} == ref, f"image {i}"
    res = metric.compute()
    keys = [str(k) for k in g[f"{case}/keys"]]
    assert sorted(res.keys()) == keys
    for k, v in zip(keys, g[f"{case}/vals"]):
        assert float(res[k]) == pytest.approx(v, rel=1e-9, abs=1e-12), k
    # list form: materialise the reference-style lists and recompute through yh_match_lists
    preds = metric.all_predictions
    for i in range(pred.shape[0]):
        assert [d[0] for d in preds[i]] is not None
    res2 = metric.compute()
    for k, v in zip(keys, g[f"{case}/vals"]):
        assert float(res2[k]) == pytest.approx(v, rel=1e-9, abs=1e-12), f"list path {k}"


@pytest.mark.parametrize("lists", ["lazy", "eager"])
def test_map_lists_match_reference_order(golden, lists):
    g = golden("decode_nms")
    for profile in ("correlated", "ties"):
        pred = torch.from_numpy(g[f"{profile}/pred"]).to(DEV)
        metric = mAPMetric(num_classes=20, conf_threshold=0.01, lists=lists)
        metric.update(pred, torch.zeros_like(pred))
        tag = f"{profile}/c0.01"
        gk = g[f"{tag}/metrics_keep_k"]
        for i, dets in enumerate(metric.all_predictions):
            n = int((gk[i] >= 0).sum())
            assert len(dets) == n
            assert [d[0] for d in dets] == g[f"{tag}/metrics_cls"][i, :n].tolist()
            assert [d[1] for d in dets] == g[f"{tag}/metrics_conf"][i, :n].tolist()
            assert np.array([d[2] for d in dets]).reshape(n, 4).tobytes() == g[f"{tag}/metrics_box"][i, :n].tobytes()
        assert all(len(gts) == 0 for gts in metric.all_ground_truths)


def test_list_modes_agree_and_lazy_mode_guards_against_inplace_edits(golden):
    """lists="lazy" rebuilds all_predictions / all_ground_truths from the retained batch, lists="eager" from what
    the update kernel stored: same lists.  A batch edited in place after update() makes the lazy rebuild raise."""
    from yolo_v1_b200 import _lib

    g = golden("map_small")
    pred, tgt = torch.from_numpy(g["corr64/pred"]).to(DEV), torch.from_numpy(g["corr64/tgt"]).to(DEV)
    a, b = mAPMetric(num_classes=20, lists="lazy"), mAPMetric(num_classes=20, lists="eager")
    for m in (a, b):
        m.update(pred[:40], tgt[:40])
        m.update(pred[40:], tgt[40:])
    assert a._chunks[0].rec_raw is None and b._chunks[0].rec_raw is not None
    assert a.all_predictions == b.all_predictions and a.all_ground_truths == b.all_ground_truths
    assert len(a.all_predictions) == 64 and sum(len(x) for x in a.all_ground_truths) > 0
    assert a.compute() == b.compute()
    c = mAPMetric(num_classes=20)
    p2 = pred.clone()
    c.update(p2, tgt)
    p2.mul_(0.5)
    with pytest.raises(_lib.YoloHeadError):
        c.all_predictions
    d = mAPMetric(num_classes=20)
    d.update(pred.cpu(), tgt.cpu())          # host tensors: the metric's own device copy is retained
    assert d.all_predictions == a.all_predictions


@pytest.mark.parametrize("n,profile,conf_thr", [(4952, "correlated", 0.01), (3000, "uniform", 0.3),
                                                (2000, "ties", 0.01), (2000, "realistic", 0.05)])
def test_map_matches_oracle_seeded(n, profile, conf_thr):
    tgt = synth.make_targets(n, seed=51, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(n, profile, seed=50, targets=tgt)
    thr = [0.5 + 0.05 * i for i in range(10)]
    st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), conf_thr, 0.4, n_threads=8)
    ref = c_oracle.compute(st, thr, n_threads=8)
    metric = mAPMetric(num_classes=20, conf_threshold=conf_thr)
    # several batches, as evaluate_model feeds it
    for lo in range(0, n, 1024):
        metric.update(pred[lo:lo + 1024].to(DEV), tgt[lo:lo + 1024].to(DEV))
    n_keep = torch.cat([c.n_keep for c in metric._chunks]).cpu().numpy()
    np.testing.assert_array_equal(n_keep, st["n_det"])
    ks = torch.cat([c.rec_k for c in metric._chunks]).cpu().numpy()
    fl = torch.cat([c.rec_flags for c in metric._chunks]).cpu().numpy().view(np.uint64) & np.uint64((1 << 56) - 1)
    mine = _flags_by_candidate(n_keep, ks, fl)
    for i in range(n):
        want = {int(st["det_k"][i, q]): int(ref["flags"][i, q]) for q in range(st["n_det"][i])}
        assert mine[i] == want, f"image {i}"
    res = metric.compute()
    want = results.assemble(ref["ap"], ref["overall"], ref["sizes"], thr, 20, n)
    assert sorted(res) == sorted(want)
    for k in want:
        assert float(res[k]) == pytest.approx(float(want[k]), rel=1e-9, abs=1e-12), k


def test_map_big_golden(golden):
    """configs[2] size (4,952 images): the reference's own compute() dict, n_keep per image and the
    hashes of its ordered keep lists and TP flag words.  The inputs come from synth.make_exact
    (integer hash + IEEE-exact operators), so a hash mismatch is a failure, not a skip."""
    import hashlib

    def sha(*arrays):
        h = hashlib.sha256()
        for a in arrays:
            h.update(np.ascontiguousarray(a).tobytes())
        return h.hexdigest()[:16]

    g = golden("map_big")
    for name in ("exact512", "exact4952"):
        n = int(g[f"{name}/n"])
        pred, tgt = synth.make_exact(n, seed=50)
        assert sha(pred.numpy(), tgt.numpy()) == str(g[f"{name}/input_sha"]), "synth.make_exact is not bit-reproducible here"
        metric = mAPMetric(num_classes=20, conf_threshold=0.01)
        metric.update(pred.to(DEV), tgt.to(DEV))
        ch = metric._chunks[0]
        n_keep = ch.n_keep.cpu().numpy()
        np.testing.assert_array_equal(n_keep, g[f"{name}/n_keep"])
        # ordered keep lists (the reference's class-grouped order) and the TP flag words in that order,
        # as make_golden.py stores them (-1 / 0 padded)
        live = np.arange(98)[None, :] < n_keep[:, None]
        r = head.decode_nms(pred.to(DEV), 0.01, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, want_boxes=False)
        np.testing.assert_array_equal(r.n_keep.cpu().numpy(), n_keep)
        keep_k = np.where(live, r.keep_k.cpu().numpy().astype(np.int32), -1).astype(np.int32)
        assert sha(keep_k) == str(g[f"{name}/keep_k_sha"])
        by_k = np.zeros((n, 98), np.uint64)          # flag word of candidate k (records are in sorted order)
        rec_k = ch.rec_k.cpu().numpy().astype(np.int64)
        fl = ch.rec_flags.cpu().numpy().view(np.uint64) & np.uint64((1 << 40) - 1)
        rows = np.repeat(np.arange(n), 98).reshape(n, 98)
        by_k[rows[live], rec_k[live]] = fl[live]
        ref_order = np.where(live, np.take_along_axis(by_k, np.maximum(keep_k, 0).astype(np.int64), 1), np.uint64(0))
        assert sha(ref_order.astype(np.uint64)) == str(g[f"{name}/flags_sha"])
        res = metric.compute()
        keys = [str(k) for k in g[f"{name}/keys"]]
        assert sorted(res) == keys
        for k, v in zip(keys, g[f"{name}/vals"]):
            assert float(res[k]) == pytest.approx(v, rel=1e-9, abs=1e-12), k


def test_sort_stability_and_scale():
    # 65,536 images (BASELINE config 2 size): per-class records must come out ordered by
    # (confidence desc, image, per-image order) and AP must not depend on how the batch is chunked
    n = 65536
    tgt = synth.make_targets(n, seed=91, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(n, "correlated", seed=90, targets=tgt)
    m1 = mAPMetric(num_classes=20, conf_threshold=0.3)
    m1.update(pred.to(DEV), tgt.to(DEV))
    r1 = m1.compute()
    m2 = mAPMetric(num_classes=20, conf_threshold=0.3)
    for lo in range(0, n, 10000):
        m2.update(pred[lo:lo + 10000].to(DEV), tgt[lo:lo + 10000].to(DEV))
    r2 = m2.compute()
    assert r1 == r2
    sub = 6000
    st = c_oracle.update_batch(pred[:sub].numpy(), tgt[:sub].numpy(), 0.3, 0.4, n_threads=8)
    ref = c_oracle.compute(st, m1.iou_thresholds, n_threads=8)
    m3 = mAPMetric(num_classes=20, conf_threshold=0.3)
    m3.update(pred[:sub].to(DEV), tgt[:sub].to(DEV))
    want = results.assemble(ref["ap"], ref["overall"], ref["sizes"], m1.iou_thresholds, 20, sub)
    got = m3.compute()
    for k in want:
        assert float(got[k]) == pytest.approx(float(want[k]), rel=1e-9, abs=1e-12), k


# ----------------------------------------------------------------------------- partition (multi-GPU exchange)
@pytest.mark.gpu
@pytest.mark.parametrize("m,parts", [(0, 2), (1, 2), (1023, 3), (1024, 8), (50_000, 8), (300_001, 32)])
def test_partition_records_is_a_stable_partition(m, parts):
    from yolo_v1_b200 import distributed as D

    rng = np.random.default_rng(m + parts)
    key = rng.integers(-2**62, 2**62, size=max(m, 1), dtype=np.int64)[:m]
    cls = rng.integers(0, 20, size=m, dtype=np.int64)
    flg = (cls << 56) | np.arange(m, dtype=np.int64)
    owners = D.class_owners(20, parts)
    k, f, cnt = D.partition_records(torch.from_numpy(key).cuda(), torch.from_numpy(flg).cuda(), m, owners, parts)
    own = np.array(owners)[cls]
    perm = np.argsort(own, kind="stable")
    assert cnt.tolist() == [int((own == p).sum()) for p in range(parts)]
    assert np.array_equal(k[:m].cpu().numpy(), key[perm])
    assert np.array_equal(f[:m].cpu().numpy(), flg[perm])


# ----------------------------------------------------------------------------- target encoder
@pytest.mark.gpu
def test_encode_targets_matches_reference_golden(golden):
    from yolo_v1_b200 import targets

    g = golden("encode")
    names = [str(n) for n in g["names"]]
    boxes = [g[f"{n}/boxes"].tolist() for n in names]
    cls = [g[f"{n}/cls"].tolist() for n in names]
    out = targets.encode_targets(boxes, cls, device=DEV).cpu().numpy()
    for i, n in enumerate(names):
        assert out[i].tobytes() == g[f"{n}/target"].tobytes(), n
    one = targets.encode_target(boxes[3], cls[3], device=DEV).cpu().numpy()
    assert one.tobytes() == g[f"{names[3]}/target"].tobytes()


@pytest.mark.gpu
def test_encode_targets_matches_oracle_and_raises_like_the_reference():
    from oracle import encode_oracle
    from yolo_v1_b200 import targets

    rng = np.random.default_rng(5)
    boxes = [[[float(v) for v in rng.random(4)] for _ in range(int(k))] for k in rng.integers(0, 60, 300)]
    cls = [[int(c) for c in rng.integers(0, 20, len(b))] for b in boxes]
    out = targets.encode_targets(boxes, cls, device=DEV).cpu().numpy()
    for i in range(len(boxes)):
        assert out[i].tobytes() == encode_oracle.encode_target(boxes[i], cls[i]).tobytes(), i
    # feeds the loss / metric kernels directly
    m = mAPMetric(num_classes=20)
    t = torch.from_numpy(out[:8]).to(DEV)
    m.update(t.clone(), t)                      # a target used as its own prediction: every kept box is a TP
    res = m.compute()
    assert float(res["precision"]) == pytest.approx(1.0, abs=1e-5) and float(res["recall"]) > 0.3
    for bad_boxes, bad_cls in ([[[0.5, -1.5, 0.1, 0.1]]], [[0]]), ([[[float("nan"), 0.5, 0.1, 0.1]]], [[0]]), \
            ([[[0.5, 0.5, 0.1, 0.1]]], [[20]]):
        with pytest.raises((IndexError, ValueError)):
            encode_oracle.encode_target(bad_boxes[0], bad_cls[0])
        with pytest.raises((IndexError, ValueError)):
            targets.encode_targets(bad_boxes, bad_cls, device=DEV)
