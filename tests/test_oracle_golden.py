"""Pins the CPU oracle (oracle/) against golden vectors produced by the unmodified
reference (tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, loss_oracle, results

LOSS_CASES = ["uniform1", "uniform2", "uniform16", "uniform64", "noobj2", "bothslots8", "slot1only4",
              "ioutie4", "rawhead8", "ties8", "correlated16", "clamp4"]


@pytest.mark.parametrize("case", LOSS_CASES)
def test_loss_oracle_matches_reference(golden, case):
    g = golden("loss")
    pred, tgt = torch.from_numpy(g[f"{case}/pred"]), torch.from_numpy(g[f"{case}/tgt"])
    parts, grad = loss_oracle.yolo_loss_and_grad(pred, tgt)
    np.testing.assert_allclose(parts.double().numpy(), g[f"{case}/parts"], rtol=2e-6, atol=1e-9)
    ref = g[f"{case}/grad"]
    np.testing.assert_allclose(grad.numpy(), ref, rtol=1e-5, atol=1e-7 * max(1.0, np.abs(ref).max()))


PROFILES = ["uniform", "realistic", "ties", "rawhead", "correlated"]


@pytest.mark.parametrize("profile", PROFILES)
@pytest.mark.parametrize("conf_thr", [0.01, 0.3, 0.5])
def test_decode_nms_oracle_matches_reference(golden, profile, conf_thr):
    g = golden("decode_nms")
    pred = g[f"{profile}/pred"]
    tag = f"{profile}/c{conf_thr}"
    st = c_oracle.update_batch(pred, None, conf_thr, 0.4, c_oracle.IOU_METRICS, c_oracle.ORDER_CLASS_GROUPED)
    gk = g[f"{tag}/metrics_keep_k"]
    n_keep = (gk >= 0).sum(1)
    np.testing.assert_array_equal(st["n_det"], n_keep)
    for i in range(pred.shape[0]):
        n = n_keep[i]
        np.testing.assert_array_equal(st["det_k"][i, :n], gk[i, :n])
        np.testing.assert_array_equal(st["det_cls"][i, :n], g[f"{tag}/metrics_cls"][i, :n])
        assert st["det_conf"][i, :n].tobytes() == g[f"{tag}/metrics_conf"][i, :n].tobytes()
        assert st["det_box"][i, :n].tobytes() == g[f"{tag}/metrics_box"][i, :n].tobytes()
    si = c_oracle.update_batch(pred, None, conf_thr, 0.4, c_oracle.IOU_INFER, c_oracle.ORDER_GLOBAL_CONF)
    ik, raises = g[f"{tag}/infer_keep_k"], g[f"{tag}/infer_raises"]
    for i in range(pred.shape[0]):
        if raises[i]:
            # the reference raised a pydantic ValidationError: some decoded candidate above
            # the threshold has a box / confidence outside [0,1]
            k, cls, conf, box = c_oracle.decode(pred[i], conf_thr)
            assert ((box < 0) | (box > 1)).any() or (conf > 1).any() or (conf < 0).any()
            continue
        n = (ik[i] >= 0).sum()
        assert si["n_det"][i] == n
        np.testing.assert_array_equal(si["det_k"][i, :n], ik[i, :n])


MAP_CASES = ["corr64", "ties48", "uniform32", "rawhead24", "corr40_thr3", "dupties48"]


@pytest.mark.parametrize("case", MAP_CASES)
def test_map_oracle_matches_reference(golden, case):
    g = golden("map_small")
    pred, tgt = g[f"{case}/pred"], g[f"{case}/tgt"]
    thr = [float(t) for t in g[f"{case}/thresholds"]]
    st = c_oracle.update_batch(pred, tgt, float(g[f"{case}/conf_thr"]), 0.4)
    gk = g[f"{case}/keep_k"]
    np.testing.assert_array_equal(st["n_det"], (gk >= 0).sum(1))
    for i in range(pred.shape[0]):
        n = st["n_det"][i]
        np.testing.assert_array_equal(st["det_k"][i, :n], gk[i, :n])
    out = c_oracle.compute(st, thr, n_threads=4)
    T = len(thr)
    ap_bits = np.uint64((1 << (4 * T)) - 1)
    np.testing.assert_array_equal(out["flags"] & ap_bits, g[f"{case}/flags"])
    res = results.assemble(out["ap"], out["overall"], out["sizes"], thr, 20, pred.shape[0])
    keys = [str(k) for k in g[f"{case}/keys"]]
    assert sorted(res.keys()) == keys
    for k, v in zip(keys, g[f"{case}/vals"]):
        assert float(res[k]) == pytest.approx(v, rel=1e-12, abs=1e-15), k
    # overall pass TP bit == (all-GT subset, IoU 0.5) TP bit whenever 0.5 is a threshold
    if 0.5 in thr:
        t5 = thr.index(0.5)
        a = (out["flags"] >> np.uint64(4 * T)) & np.uint64(1)
        b = (out["flags"] >> np.uint64(t5)) & np.uint64(1)
        np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("name", ["exact512", "exact4952"])
def test_map_oracle_matches_reference_at_config3_size(golden, name):
    """tests/golden/map_big.npz: the reference's compute() on 512 / 4,952 regenerated images."""
    import hashlib

    from yolo_v1_b200 import synth

    def sha(*arrays):
        h = hashlib.sha256()
        for a in arrays:
            h.update(np.ascontiguousarray(a).tobytes())
        return h.hexdigest()[:16]

    g = golden("map_big")
    n = int(g[f"{name}/n"])
    pred, tgt = synth.make_exact(n, seed=50)
    assert sha(pred.numpy(), tgt.numpy()) == str(g[f"{name}/input_sha"]), "synth.make_exact is not bit-reproducible here"
    thr = [0.5 + 0.05 * i for i in range(10)]
    st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), 0.01, 0.4, n_threads=4)
    np.testing.assert_array_equal(st["n_det"], g[f"{name}/n_keep"])
    live = np.arange(98)[None, :] < st["n_det"][:, None]
    assert sha(np.where(live, st["det_k"], -1).astype(np.int32)) == str(g[f"{name}/keep_k_sha"])
    out = c_oracle.compute(st, thr, n_threads=4)
    assert sha((out["flags"] & np.uint64((1 << 40) - 1)).astype(np.uint64)) == str(g[f"{name}/flags_sha"])
    res = results.assemble(out["ap"], out["overall"], out["sizes"], thr, 20, n)
    keys = [str(k) for k in g[f"{name}/keys"]]
    assert sorted(res.keys()) == keys
    for k, v in zip(keys, g[f"{name}/vals"]):
        assert float(res[k]) == pytest.approx(v, rel=1e-12, abs=1e-15), k


def test_iou_known_answers():
    # tests/test_metrics.py:35-55,208-222 and tests/test_yolo.py:196-227 of the reference
    assert c_oracle.iou((0.5, 0.5, 0.2, 0.2), (0.5, 0.5, 0.2, 0.2)) == pytest.approx(1.0, abs=1e-5)
    assert c_oracle.iou((0.2, 0.2, 0.1, 0.1), (0.8, 0.8, 0.1, 0.1)) == 0.0
    assert 0 < c_oracle.iou((0.5, 0.5, 0.4, 0.4), (0.6, 0.6, 0.4, 0.4)) < 1
    assert c_oracle.iou((0.5, 0.5, 0.0, 0.0), (0.5, 0.5, 0.2, 0.2)) == 0.0
    assert c_oracle.iou((0.5, 0.5, 0.0, 0.0), (0.5, 0.5, 0.0, 0.0)) == 0.0
    assert c_oracle.iou((0.5, 0.5, 0.3, 0.3), (0.5, 0.5, 0.3, 0.3), c_oracle.IOU_INFER) == pytest.approx(1.0, abs=1e-4)


def test_nms_known_answer():
    # tests/test_metrics.py:98-117 of the reference: keeps 2 of 3
    keep = c_oracle.nms_lists([0, 0, 1], [0.9, 0.8, 0.85],
                              [(0.5, 0.5, 0.2, 0.2), (0.52, 0.52, 0.2, 0.2), (0.7, 0.7, 0.15, 0.15)], 0.5)
    assert list(keep) == [0, 2]


def test_encode_oracle_matches_reference(golden):
    from oracle import encode_oracle

    g = golden("encode")
    for name in [str(n) for n in g["names"]]:
        got = encode_oracle.encode_target(g[f"{name}/boxes"].tolist(), g[f"{name}/cls"].tolist())
        assert got.tobytes() == g[f"{name}/target"].tobytes(), name


@pytest.mark.parametrize("name", ["m64", "m200", "m333"])
def test_head_linear_oracle_matches_reference(golden, name):
    """tests/golden/head_linear.npz (the reference's nn.Linear + view + YOLOLoss + backward): the checker smoke() uses for
    the fused kernel -- an fp32 CPU Linear over synth.head_linear_inputs followed by loss_oracle -- reproduces it."""
    from yolo_v1_b200 import synth

    g = golden("head_linear")
    M, seed = int(g[f"{name}/M"]), int(g[f"{name}/seed"])
    x, w, b, tgt = (torch.from_numpy(a) for a in synth.head_linear_inputs(M, seed))
    pred = torch.nn.functional.linear(x, w, b).view(M, 7, 7, 30)
    parts, grad = loss_oracle.yolo_loss_and_grad(pred, tgt)
    np.testing.assert_allclose(parts.double().numpy(), g[f"{name}/parts"], rtol=1e-6)
    np.testing.assert_allclose(pred[:4].numpy(), g[f"{name}/pred_head"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(grad[:4].numpy(), g[f"{name}/dpred_head"], rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(grad.reshape(M, 1470).double().sum(0).numpy(), g[f"{name}/db"], rtol=1e-4, atol=1e-6)
