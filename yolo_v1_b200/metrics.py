"""mAPMetric / evaluate_model -- drop-in for the reference's metric module on CUDA.

Public surface (constructor, ``update`` / ``compute`` / ``reset``, the private
helpers its tests call, the two public list attributes, every result key) follows
/root/reference/src/yolo/metrics.py:17-171, 654-714.  The work is done by the
kernels behind include/yolohead.h:

  update()   one fused launch per batch (yh_map_update): decode, per-class NMS,
             ground-truth parse and the image-local TP/FP matching for every
             threshold and size subset.  Results stay on the device.
  compute()  yh_compact_records + yh_ap_reduce (stable radix sort, cumulative TP,
             11-point AP) and a 4*C*T-double read-back; the means over that small
             table are taken with numpy exactly where the reference takes them.

``all_predictions`` / ``all_ground_truths`` are real Python lists, built lazily
the first time they are touched (the reference's tests append to them by hand).
By default (``lists="lazy"``) update() only keeps a reference to the batch it was
given and the lists are rebuilt from it by yh_decode_nms + yh_parse_gt, so the
update kernel writes nothing but the 16-byte records; ``lists="eager"`` makes the
update kernel write the list data next to the records instead (no reference to the
caller's tensors is kept; S=7, B=2, C=20 only -- other head shapes always rebuild).  Once touched the lists are authoritative: compute()
then uploads them and runs the list-form kernels (yh_match_lists), so hand-edited
state gives the reference's answer too.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import _lib, head

EPSILON = 1e-6
_SIZES = ("small", "medium", "large")


def _cuda_device(hint=None) -> torch.device:
    if hint is not None and torch.device(hint).type == "cuda":
        return torch.device(hint)
    if not torch.cuda.is_available():
        raise _lib.YoloHeadError("no CUDA device: the YOLOv1 head kernels have no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _conf_from_key(key: np.ndarray) -> np.ndarray:
    """Inverse of the kernels' ~orderable(conf) (common.cuh) -- a bit-level reinterpretation."""
    k = ~key.astype(np.uint64)
    neg = (k >> np.uint64(63)) == 0
    bits = np.where(neg, ~k, k & np.uint64(0x7FFFFFFFFFFFFFFF))
    return bits.view(np.float64)


_CLASS_KEYS: Dict[int, tuple] = {}


def _class_keys(C: int):
    """The per-class result keys (metrics.py:131-139), built once per class count."""
    k = _CLASS_KEYS.get(C)
    if k is None:
        k = ([f"AP50_class_{c}" for c in range(C)], [f"AP75_class_{c}" for c in range(C)],
             [f"AP50:95_class_{c}" for c in range(C)])
        _CLASS_KEYS[C] = k
    return k


class mAPMetric:
    """Mean average precision (PASCAL VOC 11-point) at one or more IoU thresholds."""

    def __init__(self, num_classes: int, iou_thresholds: List[float] = None, conf_threshold: float = 0.01,
                 nms_threshold: float = 0.4, S: int = 7, B: int = 2, *, lists: str = "lazy"):
        if lists not in ("lazy", "eager"):
            raise ValueError("lists must be 'lazy' or 'eager'")
        # extension (keyword-only): where all_predictions / all_ground_truths come from, see the module docstring
        self.lists = lists
        self.num_classes = num_classes
        if iou_thresholds is None:
            self.iou_thresholds = [0.5 + 0.05 * i for i in range(10)]
        elif isinstance(iou_thresholds, (int, float)):
            self.iou_thresholds = [float(iou_thresholds)]
        else:
            self.iou_thresholds = list(iou_thresholds)
        self.conf_threshold = conf_threshold
        self.nms_threshold = nms_threshold
        self.S = S
        self.B = B
        self._clear()

    # ------------------------------------------------------------------ state
    def _clear(self):
        self._pred_lists: list = []
        self._gt_lists: list = []
        self._chunks: list[head.UpdateChunk] = []   # device batches not yet turned into lists
        self._gt_count = None                        # device int32 [4*C], fused path only
        self._lists_live = False                     # True once the Python lists were handed out
        self._device = None

    def reset(self):
        """Forget everything accumulated so far."""
        self._clear()

    def _materialise(self):
        """Turn pending device batches into the reference's list-of-tuples form."""
        for ch in self._chunks:
            if ch.src is not None:
                dets, gts = self._lists_from_batch(ch)
            else:
                dets, gts = self._lists_from_records(ch)
            self._pred_lists.extend(dets)
            self._gt_lists.extend(gts)
        self._chunks = []
        self._gt_count = None

    def _lists_from_batch(self, ch):
        """lists="lazy": decode + NMS (reference order) and ground-truth parse of the retained batch."""
        p, t, vp, vt = ch.src
        if p._version != vp or t._version != vt:
            raise _lib.YoloHeadError(
                "all_predictions / all_ground_truths: a tensor passed to update() was modified in place afterwards, "
                "so the lists can no longer be rebuilt from it; construct mAPMetric(..., lists='eager') to have "
                "update() store the list data itself")
        S, B, C = ch.shape
        r = head.decode_nms(p, self.conf_threshold, self.nms_threshold, head.IOU_METRICS, head.ORDER_CLASS_GROUPED,
                            S, B, C, want_boxes=True)
        n_gt, gt_cls, gt_box = head.parse_gt(t, S, B, C)
        n_keep = r.n_keep.cpu().tolist()
        cls, conf, box = r.keep_cls.cpu().numpy(), r.keep_conf.cpu().numpy(), r.keep_box.cpu().numpy()
        n_gt, gcls, gbox = n_gt.cpu().tolist(), gt_cls.cpu().numpy(), gt_box.cpu().numpy()
        dets = [[(int(cls[i, q]), float(conf[i, q]), tuple(float(v) for v in box[i, q])) for q in range(n_keep[i])]
                for i in range(ch.n)]
        gts = [[(int(gcls[i, g]), tuple(float(v) for v in gbox[i, g])) for g in range(n_gt[i])] for i in range(ch.n)]
        return dets, gts

    def _lists_from_records(self, ch):
        """lists="eager": the records and side arrays the update kernel wrote (yh_list_form)."""
        n_keep = ch.n_keep.cpu().numpy()
        n_gt = ch.n_gt.cpu().numpy()
        conf = _conf_from_key(ch.rec_key.cpu().numpy())
        cls = (ch.rec_flags.cpu().numpy().view(np.uint64) >> np.uint64(56)).astype(np.int64)
        rec_pos, rec_box, gt_box = head.list_form(ch)
        pos = rec_pos.cpu().numpy()
        box = rec_box.cpu().numpy()
        gcls = ch.gt_cls.cpu().numpy()
        gbox = gt_box.cpu().numpy()
        all_dets, all_gts = [], []
        for i in range(ch.n):
            k = int(n_keep[i])
            dets = [None] * k
            for q in range(k):
                dets[int(pos[i, q])] = (int(cls[i, q]), float(conf[i, q]), tuple(float(v) for v in box[i, q]))
            all_dets.append(dets)
            all_gts.append([(int(gcls[i, g]), tuple(float(v) for v in gbox[i, g])) for g in range(int(n_gt[i]))])
        return all_dets, all_gts

    @property
    def all_predictions(self) -> list:
        self._materialise()
        self._lists_live = True
        return self._pred_lists

    @all_predictions.setter
    def all_predictions(self, value):
        self._materialise()
        self._lists_live = True
        self._pred_lists = value

    @property
    def all_ground_truths(self) -> list:
        self._materialise()
        self._lists_live = True
        return self._gt_lists

    @all_ground_truths.setter
    def all_ground_truths(self, value):
        self._materialise()
        self._lists_live = True
        self._gt_lists = value

    # ------------------------------------------------------------------ update
    def update(self, predictions: torch.Tensor, targets: torch.Tensor):
        """Accumulate a batch: predictions and targets of shape (batch, S, S, B*5 + C)."""
        # the kernels run on CUDA only; host tensors (the reference's own tests pass them) are
        # copied to the GPU first -- with no CUDA device this raises, there is no CPU path
        if not predictions.is_cuda:
            predictions = predictions.to(_cuda_device(self._device))
        dev = predictions.device
        targets = targets.to(dev)
        if self._gt_count is None or self._device != dev:
            if self._chunks and self._device != dev:
                self._materialise()
                self._lists_live = True
            self._gt_count = torch.zeros(4 * self.num_classes, dtype=torch.int32, device=dev)
            self._device = dev
        chunk = head.map_update(predictions, targets, self.conf_threshold, self.nms_threshold,
                                self.iou_thresholds, self._gt_count, self.S, self.B, self.num_classes,
                                want_lists=self.lists == "eager" and (self.S, self.B, self.num_classes) == (7, 2, 20))
        self._chunks.append(chunk)
        if self._lists_live:
            self._materialise()

    def update_with_loss(self, predictions: torch.Tensor, targets: torch.Tensor, criterion, backward: bool = True,
                         tail_stream=None):
        """Extension: update() and `criterion(predictions, targets)` (+ backward) of the same batch in ONE pass over
        the two tensors (yh_map_update_loss) -- what a validation step that wants both the loss and the metric
        does.  Returns (parts, grad): the five loss values as a (5,) fp32 device tensor in the order of the
        reference's loss_dict (total, coord, conf_obj, conf_noobj, class; loss.py:164-170) and d total / d
        predictions (None without `backward`).  Needs S=7, B=2, C=20, fp32 CUDA predictions; anything else raises
        and the caller runs the two calls separately.  tail_stream: see head.map_update_loss."""
        if (self.S, self.B, self.num_classes) != (7, 2, 20) or (criterion.S, criterion.B, criterion.C) != (7, 2, 20):
            raise NotImplementedError("update_with_loss: the fused kernel is built for S=7, B=2, C=20")
        _lib.require_cuda(predictions, "predictions")
        if predictions.dtype is not torch.float32:
            raise NotImplementedError("update_with_loss: fp32 predictions only (the loss alone takes fp16 / bf16)")
        dev = predictions.device
        targets = targets.to(dev)
        if self._gt_count is None or self._device != dev:
            if self._chunks and self._device != dev:
                self._materialise()
                self._lists_live = True
            self._gt_count = torch.zeros(4 * self.num_classes, dtype=torch.int32, device=dev)
            self._device = dev
        chunk, parts, grad = head.map_update_loss(predictions, targets, self.conf_threshold, self.nms_threshold,
                                                  self.iou_thresholds, self._gt_count, criterion.lambda_coord,
                                                  criterion.lambda_noobj, want_grad=backward, tail_stream=tail_stream)
        self._chunks.append(chunk)
        if self._lists_live:
            self._materialise()
        return parts, grad

    # ------------------------------------------------------------------ compute
    def compute(self) -> Dict[str, float]:
        """mAP50:95 / mAP50 / mAP75, per-class APs, overall precision / recall, size-bucketed mAPs."""
        n_images = len(self._pred_lists) + sum(c.n for c in self._chunks)
        if n_images == 0:
            return {"mAP50:95": 0.0, "mAP50": 0.0, "mAP75": 0.0, "precision": 0.0, "recall": 0.0}
        T, C = len(self.iou_thresholds), self.num_classes
        if self._lists_live or self._pred_lists:
            self._materialise()
            ap, tp_total, n_pred, gt_count, n_gt_total = self._reduce_lists()
        else:
            ap, tp_total, n_pred, gt_count = self._reduce_chunks()
            n_gt_total = int(gt_count.reshape(4, C)[0].sum())
        return self._results(ap, tp_total, n_pred, gt_count.reshape(4, C), n_gt_total, T, C)

    def _enqueue_reduce(self) -> torch.Tensor:
        """Device records -> packed AP result, all enqueued, nothing read back (capturable in a CUDA
        graph): the batches are compacted into one buffer behind a device-side running count, and
        yh_ap_reduce takes that count from the device."""
        T, C = len(self.iou_thresholds), self.num_classes
        dev = self._device
        cap = sum(ch.rec_key.numel() for ch in self._chunks)
        key = torch.empty(cap, dtype=torch.int64, device=dev)
        flg = torch.empty(cap, dtype=torch.int64, device=dev)
        running = torch.zeros(1, dtype=torch.int64, device=dev)
        for ch in self._chunks:
            head.compact_records(ch, key, flg, running)
        return head.ap_reduce(key, flg, cap, self._gt_count, C, T, m_dev=running)[2]

    def _results_from_packed(self, packed: torch.Tensor) -> Dict[str, float]:
        """The one device->host copy of compute() and the result dict."""
        T, C = len(self.iou_thresholds), self.num_classes
        ap, tp_total, n_pred, gt_count = head.unpack_ap(packed, C, T)
        return self._results(ap, tp_total, n_pred, gt_count, int(gt_count[0].sum()), T, C)

    def _reduce_chunks(self):
        T, C = len(self.iou_thresholds), self.num_classes
        ap, tp_total, n_pred, gt_count = head.unpack_ap(self._enqueue_reduce(), C, T)
        return ap, tp_total, n_pred, gt_count.reshape(-1)

    def _reduce_lists(self):
        T, C = len(self.iou_thresholds), self.num_classes
        dev = _cuda_device(self._device)
        preds, gts = self._pred_lists, self._gt_lists
        n = min(len(preds), len(gts))  # the reference zips the two lists
        det_off = np.zeros(n + 1, np.int64)
        gt_off = np.zeros(n + 1, np.int64)
        for i in range(n):
            det_off[i + 1] = det_off[i] + len(preds[i])
            gt_off[i + 1] = gt_off[i] + len(gts[i])
        M, G = int(det_off[-1]), int(gt_off[-1])
        det_cls = np.fromiter((d[0] for p in preds[:n] for d in p), np.int32, M)
        det_conf = np.fromiter((d[1] for p in preds[:n] for d in p), np.float64, M)
        det_box = np.array([d[2] for p in preds[:n] for d in p], np.float64).reshape(M, 4)
        gt_cls = np.fromiter((g[0] for gl in gts[:n] for g in gl), np.int32, G)
        gt_box = np.array([g[1] for gl in gts[:n] for g in gl], np.float64).reshape(G, 4)
        to = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
        pad = lambda a: a if a.size else np.zeros((1,) + a.shape[1:], a.dtype)  # noqa: E731
        gt_count = torch.zeros(4 * C, dtype=torch.int32, device=dev)
        key, flg = head.match_lists(to(det_off), to(pad(det_cls)), to(pad(det_conf)), to(pad(det_box)),
                                    to(gt_off), to(pad(gt_cls)), to(pad(gt_box)), C, self.iou_thresholds, gt_count)
        _, _, packed = head.ap_reduce(key, flg, M, gt_count, C, T)
        ap, tp_total, n_pred, gtc = head.unpack_ap(packed, C, T)
        return ap, tp_total, n_pred, gtc.reshape(-1), G

    def _results(self, ap, tp_total, n_pred, gt_count, n_gt_total, T, C) -> Dict[str, float]:  # noqa: C901
        """Result dict with the reference's keys and its np.mean call sites (metrics.py:126-171, 531-564)."""
        self._last_raw = (ap, tp_total, n_pred, gt_count)      # what result_digest() hashes
        thr = self.iou_thresholds
        if len(set(thr)) == len(thr):
            return self._results_fast(ap, tp_total, n_pred, gt_count, n_gt_total, T, C)
        out: Dict[str, float] = {}
        by_thr = {t: [] for t in thr}
        for c in range(C):
            per_class = [float(ap[0, c, ti]) for ti in range(T)]
            for ti, t in enumerate(thr):
                by_thr[t].append(per_class[ti])
                if t == 0.5:
                    out[f"AP50_class_{c}"] = per_class[ti]
                elif t == 0.75:
                    out[f"AP75_class_{c}"] = per_class[ti]
            out[f"AP50:95_class_{c}"] = np.mean(per_class)
        if 0.5 in thr:
            out["mAP50"] = np.mean(by_thr[0.5])
        if 0.75 in thr:
            out["mAP75"] = np.mean(by_thr[0.75])
        out["mAP50:95"] = np.mean([v for vals in by_thr.values() for v in vals])
        fp_total = n_pred - tp_total
        out["precision"] = tp_total / (tp_total + fp_total + EPSILON)
        out["recall"] = tp_total / (n_gt_total + EPSILON)
        counts = [int(gt_count[s].sum()) for s in (1, 2, 3)]
        for s, name in enumerate(_SIZES, start=1):
            if counts[s - 1] == 0:
                out[f"mAP50:95_{name}"] = 0.0
                out[f"mAP50_{name}"] = 0.0
                out[f"mAP75_{name}"] = 0.0
                continue
            by_thr = {t: [] for t in thr}  # a repeated threshold appends twice to one list, as metrics.py:543-550
            for c in range(C):
                for ti, t in enumerate(thr):
                    by_thr[t].append(float(ap[s, c, ti]))
            if 0.5 in thr:
                out[f"mAP50_{name}"] = np.mean(by_thr[0.5])
            if 0.75 in thr:
                out[f"mAP75_{name}"] = np.mean(by_thr[0.75])
            out[f"mAP50:95_{name}"] = np.mean([v for vals in by_thr.values() for v in vals])
        for name, cnt in zip(_SIZES, counts):
            out[f"num_{name}_objects"] = cnt
        return out

    def _results_fast(self, ap, tp_total, n_pred, gt_count, n_gt_total, T, C) -> Dict[str, float]:
        """_results for distinct thresholds, a dozen numpy calls instead of ~35.

        np.mean(list) is add.reduce over the contiguous float64 copy of the list divided by its
        length; the row / column / flattened reductions below run the same pairwise summation
        on the same element order, so every value is bit-identical to the loop above."""
        thr = self.iou_thresholds
        ap = np.ascontiguousarray(ap, dtype=np.float64)
        out: Dict[str, float] = {}
        i50 = thr.index(0.5) if 0.5 in thr else -1
        i75 = thr.index(0.75) if 0.75 in thr else -1
        a0 = ap[0]
        cls_mean = np.add.reduce(a0, axis=1) / T
        col50 = a0[:, i50].tolist() if i50 >= 0 else None
        col75 = a0[:, i75].tolist() if i75 >= 0 else None
        k50, k75, kall = _class_keys(C)
        per_class = list(cls_mean)           # np.float64 items, as np.mean returns
        if i50 >= 0 and i75 >= 0:            # same insertion order as the loop above
            for c in range(C):
                out[k50[c]] = col50[c]
                out[k75[c]] = col75[c]
                out[kall[c]] = per_class[c]
        else:
            for c in range(C):
                if i50 >= 0:
                    out[k50[c]] = col50[c]
                if i75 >= 0:
                    out[k75[c]] = col75[c]
                out[kall[c]] = per_class[c]

        def means(a, prefix, suffix):
            if i50 >= 0:
                out[f"{prefix}50{suffix}"] = np.add.reduce(np.ascontiguousarray(a[:, i50])) / C
            if i75 >= 0:
                out[f"{prefix}75{suffix}"] = np.add.reduce(np.ascontiguousarray(a[:, i75])) / C
            out[f"{prefix}50:95{suffix}"] = np.add.reduce(np.ascontiguousarray(a.T).reshape(-1)) / (C * T)

        means(a0, "mAP", "")
        fp_total = n_pred - tp_total
        out["precision"] = tp_total / (tp_total + fp_total + EPSILON)
        out["recall"] = tp_total / (n_gt_total + EPSILON)
        counts = [int(gt_count[s].sum()) for s in (1, 2, 3)]
        for s, name in enumerate(_SIZES, start=1):
            if counts[s - 1] == 0:
                out[f"mAP50:95_{name}"] = 0.0
                out[f"mAP50_{name}"] = 0.0
                out[f"mAP75_{name}"] = 0.0
            else:
                means(ap[s], "mAP", f"_{name}")
        for name, cnt in zip(_SIZES, counts):
            out[f"num_{name}_objects"] = cnt
        return out

    def result_digest(self) -> str:
        """sha256 over everything the last compute() read back: the full 4 x C x T AP table (every size subset and
        threshold, not only the entries that reach the result dict), the TP total, the record count and the
        ground-truth counts.  Two runs that order, exchange or flag a single record differently disagree here."""
        import hashlib

        ap, tp_total, n_pred, gt_count = self._last_raw
        h = hashlib.sha256()
        h.update(np.ascontiguousarray(ap, dtype=np.float64).tobytes())
        h.update(np.asarray([tp_total, n_pred], dtype=np.int64).tobytes())
        h.update(np.ascontiguousarray(gt_count, dtype=np.int64).tobytes())
        return h.hexdigest()[:16]

    # --------------------------------------------------- helpers the reference exposes
    def _parse_predictions(self, pred: torch.Tensor) -> List[Tuple[int, float, Tuple[float, float, float, float]]]:
        """(S,S,B*5+C) tensor -> [(class_id, confidence, (x, y, w, h))] in cell/box order (metrics.py:173-218)."""
        if not pred.is_cuda:
            pred = pred.to(_cuda_device(self._device))
        r = head.decode_nms(pred, self.conf_threshold, float("nan"), head.IOU_METRICS, head.ORDER_CLASS_GROUPED,
                            self.S, self.B, self.num_classes)
        n = int(r.n_keep[0].item())
        k = r.keep_k[0, :n].cpu().tolist()
        cls = r.keep_cls[0, :n].cpu().tolist()
        conf = r.keep_conf[0, :n].cpu().tolist()
        box = r.keep_box[0, :n].cpu().tolist()
        order = sorted(range(n), key=k.__getitem__)
        return [(int(cls[i]), float(conf[i]), tuple(box[i])) for i in order]

    def _parse_ground_truth(self, target: torch.Tensor) -> List[Tuple[int, Tuple[float, float, float, float]]]:
        """(S,S,B*5+C) target -> [(class_id, (x, y, w, h))] (metrics.py:220-256)."""
        if not target.is_cuda:
            target = target.to(_cuda_device(self._device))
        n_gt, gt_cls, gt_box = head.parse_gt(target, self.S, self.B, self.num_classes)
        n = int(n_gt.item())
        return [(int(c), tuple(b)) for c, b in zip(gt_cls[0, :n].cpu().tolist(), gt_box[0, :n].cpu().tolist())]

    def _apply_nms(self, detections):
        """Per-class greedy NMS over a list of (class_id, confidence, bbox) tuples (metrics.py:258-296)."""
        if len(detections) == 0:
            return []
        dev = _cuda_device(self._device)
        n_keep, keep_idx = head.nms_lists([0, len(detections)], [d[0] for d in detections],
                                          [d[1] for d in detections], [d[2] for d in detections],
                                          self.nms_threshold, head.IOU_METRICS, head.ORDER_CLASS_GROUPED, dev)
        n = int(n_keep[0].item())
        return [detections[i] for i in keep_idx[:n].cpu().tolist()]

    def _calculate_iou(self, box1, box2) -> float:
        """IoU of two (x_center, y_center, w, h) boxes; 0.0 when the union is empty (metrics.py:298-341)."""
        dev = _cuda_device(self._device)
        return float(head.iou_pairs([box1], [box2], head.IOU_METRICS, dev).item())


def evaluate_model(model: nn.Module, dataloader, device: str, num_classes: int = 20,
                   iou_thresholds: List[float] = None, conf_threshold: float = 0.01,
                   nms_threshold: float = 0.4, S: int = 7, B: int = 2) -> Dict[str, float]:
    """Run `model` over `dataloader` and return mAPMetric.compute() (metrics.py:654-714).

    Predictions go from the model straight into the fused update kernel; nothing is
    copied to the host until compute()."""
    from tqdm import tqdm

    model.eval()
    metric = mAPMetric(num_classes=num_classes, iou_thresholds=iou_thresholds, conf_threshold=conf_threshold,
                       nms_threshold=nms_threshold, S=S, B=B)
    with torch.no_grad():
        for images, targets in tqdm(dataloader, desc="Evaluating", unit="batch"):
            images = images.to(device, non_blocking=True)
            targets = targets.to(device, non_blocking=True)
            metric.update(model(images), targets)
    return metric.compute()
