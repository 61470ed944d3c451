"""ctypes front end of oracle/liboracle.so (the C restatement in yolo_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(yolo_v1_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")

IOU_METRICS, IOU_INFER = 0, 1
ORDER_CLASS_GROUPED, ORDER_GLOBAL_CONF = 0, 1


def build(force: bool = False) -> str:
    """Compile liboracle.so with the recipe in oracle/Makefile."""
    src = os.path.join(_HERE, "yolo_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.yo_iou.restype = C.c_double
        _lib.yo_iou.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def iou(box1, box2, mode=IOU_METRICS) -> float:
    a = np.asarray(box1, dtype=np.float64)
    b = np.asarray(box2, dtype=np.float64)
    return float(lib().yo_iou(_p(a), _p(b), int(mode)))


def decode(pred: np.ndarray, conf_thr: float, S=7, B=2, Cn=20):
    """One image -> (k, cls, conf, box[n,4]) in emission order."""
    pred = np.ascontiguousarray(pred, dtype=np.float32)
    K = S * S * B
    k = np.zeros(K, np.int32); cls = np.zeros(K, np.int32)
    conf = np.zeros(K, np.float64); box = np.zeros((K, 4), np.float64)
    n = lib().yo_decode(_p(pred), S, B, Cn, C.c_double(conf_thr), _p(k), _p(cls), _p(conf), _p(box))
    return k[:n], cls[:n], conf[:n], box[:n]


def parse_gt(tgt: np.ndarray, S=7, B=2, Cn=20):
    tgt = np.ascontiguousarray(tgt, dtype=np.float32)
    G = S * S
    cell = np.zeros(G, np.int32); cls = np.zeros(G, np.int32); box = np.zeros((G, 4), np.float64)
    n = lib().yo_parse_gt(_p(tgt), S, B, Cn, _p(cell), _p(cls), _p(box))
    return cell[:n], cls[:n], box[:n]


def nms_lists(cls, conf, box, thr, iou_mode=IOU_METRICS, order=ORDER_CLASS_GROUPED):
    """Generic greedy NMS; returns indices into the input list, reference order."""
    cls = np.ascontiguousarray(cls, dtype=np.int32)
    conf = np.ascontiguousarray(conf, dtype=np.float64)
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(-1, 4)
    n = len(cls)
    keep = np.zeros(max(n, 1), np.int32)
    nk = lib().yo_nms_lists(n, _p(cls), _p(conf), _p(box), C.c_double(thr), int(iou_mode),
                            int(order), _p(keep))
    return keep[:nk].copy()


def update_batch(pred, tgt, conf_thr, nms_thr, iou_mode=IOU_METRICS, order=ORDER_CLASS_GROUPED,
                 S=7, B=2, Cn=20, n_threads=1):
    """decode + GT parse + NMS for a batch.  Returns a dict of fixed-stride arrays."""
    pred = np.ascontiguousarray(pred, dtype=np.float32)
    N = pred.shape[0]
    K, G = S * S * B, S * S
    out = dict(
        N=N, K=K, G=G,
        n_det=np.zeros(N, np.int32), det_k=np.zeros((N, K), np.int32),
        det_cls=np.zeros((N, K), np.int32), det_conf=np.zeros((N, K), np.float64),
        det_box=np.zeros((N, K, 4), np.float64),
        n_gt=np.zeros(N, np.int32), gt_cls=np.zeros((N, G), np.int32),
        gt_box=np.zeros((N, G, 4), np.float64),
    )
    tp = None
    if tgt is not None:
        tgt = np.ascontiguousarray(tgt, dtype=np.float32)
        tp = _p(tgt)
    rc = lib().yo_update_batch(_p(pred), tp, N, S, B, Cn, C.c_double(conf_thr), C.c_double(nms_thr),
                               int(iou_mode), int(order), int(n_threads),
                               _p(out["n_det"]), _p(out["det_k"]), _p(out["det_cls"]),
                               _p(out["det_conf"]), _p(out["det_box"]),
                               _p(out["n_gt"]), _p(out["gt_cls"]), _p(out["gt_box"]))
    if rc != 0:
        raise RuntimeError(f"yo_update_batch failed: {rc}")
    return out


def compute(state, thresholds, Cn=20, n_threads=1):
    """AP table / TP flags / counts from a fixed-stride state (see update_batch)."""
    thr = np.ascontiguousarray(thresholds, dtype=np.float64)
    T = len(thr)
    N, K, G = state["N"], state["K"], state["G"]
    ap = np.zeros((4, Cn, T), np.float64)
    flags = np.zeros((N, K), np.uint64)
    overall = np.zeros(3, np.int64)
    sizes = np.zeros(3, np.int64)
    rc = lib().yo_compute(N, K, G, Cn, _p(state["n_det"]), _p(state["det_cls"]), _p(state["det_conf"]),
                          _p(state["det_box"]), _p(state["n_gt"]), _p(state["gt_cls"]),
                          _p(state["gt_box"]), _p(thr), T, int(n_threads), _p(ap), _p(flags),
                          _p(overall), _p(sizes))
    if rc != 0:
        raise RuntimeError(f"yo_compute failed: {rc}")
    return dict(ap=ap, flags=flags, overall=overall, sizes=sizes)
