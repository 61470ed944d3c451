"""Head shapes other than S=7, B=2, C=20 (the reference's S / B / C constructor arguments: loss.py:28-53,
metrics.py:43-71, inference.py:170-171).  tests/golden/generic.npz holds what the UNMODIFIED reference computes
for eight shapes (tests/golden/make_generic_golden.py); the oracle is pinned to it on the CPU, and the CUDA
path (csrc/generic.cu behind the same entry points) is held to it on the GPU: keep lists bit-exact, loss /
gradient / result dict to 1e-5 / 1e-9."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, loss_oracle, results

SHAPES = ["S7B2C3", "S7B2C1", "S7B2C32", "S4B3C5", "S11B2C20", "S3B1C2", "S5B8C7", "S7B1C20"]
PROFILES = ["uniform", "ties"]


def _shape(name):
    s, rest = name[1:].split("B")
    b, c = rest.split("C")
    return int(s), int(b), int(c)


@pytest.mark.parametrize("name", SHAPES)
@pytest.mark.parametrize("prof", PROFILES)
def test_oracle_matches_reference_on_other_shapes(golden, name, prof):
    g = golden("generic")
    S, B, C = _shape(name)
    tag = f"{name}/{prof}"
    pred, tgt = g[f"{tag}/pred"], g[f"{tag}/tgt"]
    parts, grad = loss_oracle.yolo_loss_and_grad(torch.from_numpy(pred), torch.from_numpy(tgt), S=S, B=B, C=C)
    np.testing.assert_allclose(parts.double().numpy(), g[f"{tag}/loss_parts"], rtol=3e-6, atol=1e-9)
    ref = g[f"{tag}/loss_grad"]
    np.testing.assert_allclose(grad.numpy(), ref, rtol=1e-5, atol=1e-7 * max(1.0, np.abs(ref).max()))
    st = c_oracle.update_batch(pred, tgt, 0.2, 0.4, S=S, B=B, Cn=C)
    gk = g[f"{tag}/metrics_keep_k"]
    np.testing.assert_array_equal(st["n_det"], (gk >= 0).sum(1))
    live = gk >= 0
    assert np.array_equal(np.where(live, st["det_k"], -1), gk)
    np.testing.assert_array_equal(st["n_gt"], g[f"{tag}/n_gt"])
    thr = [0.5 + 0.05 * i for i in range(10)]
    out = c_oracle.compute(st, thr, Cn=C)
    res = results.assemble(out["ap"], out["overall"], out["sizes"], thr, C, pred.shape[0])
    keys = [str(k) for k in g[f"{tag}/keys"]]
    assert sorted(res.keys()) == keys
    for k, v in zip(keys, g[f"{tag}/vals"]):
        assert float(res[k]) == pytest.approx(v, rel=1e-12, abs=1e-15), k
    si = c_oracle.update_batch(pred, None, 0.2, 0.4, c_oracle.IOU_INFER, c_oracle.ORDER_GLOBAL_CONF, S=S, B=B, Cn=C)
    ik, raises = g[f"{tag}/infer_keep_k"], g[f"{tag}/infer_raises"]
    for i in range(pred.shape[0]):
        if not raises[i]:
            n = int((ik[i] >= 0).sum())
            assert si["n_det"][i] == n and np.array_equal(si["det_k"][i, :n], ik[i, :n])


@pytest.mark.gpu
@pytest.mark.parametrize("name", SHAPES)
@pytest.mark.parametrize("prof", PROFILES)
def test_cuda_matches_reference_on_other_shapes(golden, name, prof):
    from yolo_v1_b200 import YOLOLoss, head, mAPMetric

    g = golden("generic")
    S, B, C = _shape(name)
    tag = f"{name}/{prof}"
    pred, tgt = torch.from_numpy(g[f"{tag}/pred"]), torch.from_numpy(g[f"{tag}/tgt"])
    # ---- loss + gradient (a1-a3)
    p = pred.cuda().requires_grad_(True)
    total, parts = YOLOLoss(S=S, B=B, C=C)(p, tgt.cuda())
    total.backward()
    got = np.array([parts[k] for k in ("total", "coord", "conf_obj", "conf_noobj", "class")])
    np.testing.assert_allclose(got, g[f"{tag}/loss_parts"], rtol=1e-5, atol=1e-9)
    ref = g[f"{tag}/loss_grad"]
    np.testing.assert_allclose(p.grad.cpu().numpy(), ref, rtol=1e-5, atol=1e-7 * max(1.0, np.abs(ref).max()))
    # ---- ordered keep lists, both flavours (a4-a8)
    gk = g[f"{tag}/metrics_keep_k"]
    r = head.decode_nms(pred.cuda(), 0.2, 0.4, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, S, B, C)
    nk = r.n_keep.cpu().numpy()
    np.testing.assert_array_equal(nk, (gk >= 0).sum(1))
    live = gk >= 0
    assert np.array_equal(np.where(live, r.keep_k.cpu().numpy().astype(np.int32), -1), gk)
    ik, raises = g[f"{tag}/infer_keep_k"], g[f"{tag}/infer_raises"]
    ri = head.decode_nms(pred.cuda(), 0.2, 0.4, head.IOU_INFER, head.ORDER_GLOBAL_CONF, S, B, C)
    assert np.array_equal(ri.bad_range.cpu().numpy().astype(bool), raises)
    ok = ~raises
    assert np.array_equal(np.where(ik >= 0, ri.keep_k.cpu().numpy().astype(np.int32), -1)[ok], ik[ok])
    # ---- update / compute / lists (a9-a14)
    m = mAPMetric(num_classes=C, conf_threshold=0.2, nms_threshold=0.4, S=S, B=B)
    m.update(pred[:6], tgt[:6])                      # host tensors, two batches
    m.update(pred[6:].cuda(), tgt[6:].cuda())
    res = m.compute()
    keys = [str(k) for k in g[f"{tag}/keys"]]
    assert sorted(res) == keys
    for k, v in zip(keys, g[f"{tag}/vals"]):
        assert float(res[k]) == pytest.approx(v, rel=1e-9, abs=1e-12), k
    assert [len(x) for x in m.all_ground_truths] == g[f"{tag}/n_gt"].tolist()
    for i, dets in enumerate(m.all_predictions):
        assert len(dets) == int((gk[i] >= 0).sum())
    assert m.compute() == res                        # from the materialised lists (yh_match_lists): same dict


@pytest.mark.gpu
def test_other_shapes_against_the_oracle_at_size_and_engine_shapes():
    """A few hundred images of two shapes against the C oracle (flags bit-exact), and YOLOInference taking S / B
    from the model (inference.py:170-171)."""
    import torch.nn as nn

    from yolo_v1_b200 import YOLOInference, mAPMetric, synth

    for S, B, C in ((7, 2, 3), (5, 3, 11)):
        n = 400
        tgt = synth.make_targets(n, seed=S + C, S=S, B=B, C=C, p_obj=0.1, size_mix=True)
        pred = synth.make_preds(n, "correlated", seed=S + C + 1, S=S, B=B, C=C, targets=tgt)
        thr = [0.5 + 0.05 * i for i in range(10)]
        st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), 0.05, 0.4, S=S, B=B, Cn=C, n_threads=4)
        ref = c_oracle.compute(st, thr, Cn=C, n_threads=4)
        m = mAPMetric(num_classes=C, conf_threshold=0.05, S=S, B=B)
        m.update(pred.cuda(), tgt.cuda())
        ch = m._chunks[0]
        nk = ch.n_keep.cpu().numpy()
        np.testing.assert_array_equal(nk, st["n_det"])
        ks = ch.rec_k.cpu().numpy()
        fl = ch.rec_flags.cpu().numpy().view(np.uint64) & np.uint64((1 << 56) - 1)
        for i in range(n):
            mine = {int(ks[i, q]): int(fl[i, q]) for q in range(nk[i])}
            want = {int(st["det_k"][i, q]): int(ref["flags"][i, q]) for q in range(st["n_det"][i])}
            assert mine == want, (S, B, C, i)
        got = m.compute()
        want = results.assemble(ref["ap"], ref["overall"], ref["sizes"], thr, C, n)
        for k in want:
            assert float(got[k]) == pytest.approx(float(want[k]), rel=1e-9, abs=1e-12), k

    class Net(nn.Module):
        S, B = 5, 3

        def forward(self, x):
            return x

    eng = YOLOInference(Net(), device="cuda")
    pred = torch.zeros(5, 5, 3 * 5 + 4)
    pred[1, 2, 5:10] = torch.tensor([0.5, 0.5, 0.3, 0.3, 0.9])     # box 1 of cell (1, 2)
    pred[1, 2, 15 + 2] = 0.8
    dets = eng.parse_predictions(pred, conf_threshold=0.5)
    assert len(dets) == 1 and dets[0].class_id == 2 and dets[0].bbox.x == pytest.approx((2 + 0.5) / 5)
    assert eng.predict_batch(pred[None], 0.5, 0.4)[0] == dets


@pytest.mark.gpu
def test_shapes_beyond_the_build_say_so():
    from yolo_v1_b200 import YOLOLoss, mAPMetric

    with pytest.raises(NotImplementedError):
        mAPMetric(num_classes=40).update(torch.zeros(1, 7, 7, 50).cuda(), torch.zeros(1, 7, 7, 50).cuda())   # C > 32
    with pytest.raises(NotImplementedError):
        mAPMetric(num_classes=20, S=14).update(torch.zeros(1, 14, 14, 30).cuda(), torch.zeros(1, 14, 14, 30).cuda())  # 392 candidates
    with pytest.raises(NotImplementedError):
        YOLOLoss(S=7, B=9, C=20)(torch.zeros(1, 7, 7, 65).cuda(), torch.zeros(1, 7, 7, 65).cuda())
