"""YOLOInference -- drop-in for the reference's inference engine, post-processing on CUDA.

Constructor, ``predict`` / ``parse_predictions`` / ``iou`` / ``non_max_suppression``
signatures, defaults, the DeprecationWarning for ``iou_threshold=`` and the
pydantic validation behaviour follow /root/reference/src/yolo/inference.py:12-317.
Decode and NMS run in yh_decode_nms (one warp per image); ``predict_batch`` is the
batched form the kernel is built for.
"""
from __future__ import annotations

import warnings
from collections.abc import Sequence

import torch
import torch.nn as nn

from . import _lib, head
from .schemas import BoundingBox, Detection

EPSILON = 1e-6


def _default_device() -> str:
    if torch.cuda.is_available():
        return "cuda"
    raise _lib.YoloHeadError("no CUDA device: YOLOInference post-processing has no CPU fallback")


class LazyDetections(Sequence):
    """The detections of one image as a read-only sequence: `Detection` objects (pydantic-validated, as the reference
    builds them at inference.py:197-208) are constructed when an element is first read and kept.  len(), indexing,
    slicing, iteration and == against a list of Detections behave like the eager list predict_batch() returns."""

    __slots__ = ("_cls", "_conf", "_box", "_names", "_items", "built")

    def __init__(self, cls, conf, box, class_names):
        self._cls, self._conf, self._box, self._names = cls, conf, box, class_names
        self._items: list = [None] * len(cls)
        self.built = 0                       # Detections constructed so far

    def __len__(self) -> int:
        return len(self._items)

    def _at(self, q: int) -> Detection:
        d = self._items[q]
        if d is None:
            cid = int(self._cls[q])
            b = self._box[q]
            d = Detection(class_id=cid, class_name=self._names[cid] if self._names else f"class_{cid}",
                          confidence=float(self._conf[q]),
                          bbox=BoundingBox(x=float(b[0]), y=float(b[1]), width=float(b[2]), height=float(b[3])))
            self._items[q] = d
            self.built += 1
        return d

    def __getitem__(self, q):
        if isinstance(q, slice):
            return [self._at(i) for i in range(*q.indices(len(self._items)))]
        n = len(self._items)
        if q < -n or q >= n:
            raise IndexError("LazyDetections index out of range")
        return self._at(q % n)

    def __iter__(self):
        return (self._at(i) for i in range(len(self._items)))

    def __eq__(self, other):
        if isinstance(other, (list, tuple, LazyDetections)):
            return len(other) == len(self) and all(a == b for a, b in zip(self, other))
        return NotImplemented

    def __repr__(self) -> str:
        return f"LazyDetections({len(self)} detections, {self.built} built)"

    # what a caller can ask without building anything
    @property
    def class_ids(self):
        return self._cls.astype("int64")

    @property
    def confidences(self):
        return self._conf

    @property
    def boxes(self):
        """(n, 4) float64 array of (x, y, width, height), normalised to the image."""
        return self._box


class YOLOInference:
    """Image -> list[Detection]: load, preprocess, model forward, decode, NMS."""

    def __init__(self, model: nn.Module, device: str | None = None) -> None:
        self.device = device if device is not None else _default_device()
        self.model = model.to(self.device)
        self.model.eval()
        from torchvision import transforms

        self.transform = transforms.Compose([
            transforms.Resize((448, 448)),
            transforms.ToTensor(),
            transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225]),
        ])

    # ---------------------------------------------------------------- image I/O (unchanged path)
    def load_image(self, image_path: str):
        from PIL import Image

        return Image.open(image_path).convert("RGB")

    def preprocess_image(self, image) -> torch.Tensor:
        """(1, 3, 448, 448) float32 on the model's device (inference.py:88-101).  With the model on a GPU the
        resize / scale / normalise runs there (yh_preprocess_rgb8, bit-identical to `self.transform`); a model
        kept on the host gets the reference's torchvision transform, as the backbone is its business."""
        if torch.device(self.device).type == "cuda":
            from .preprocess import preprocess_pil

            return preprocess_pil(image, self.device).unsqueeze(0)
        return self.transform(image).unsqueeze(0).to(self.device)

    # ---------------------------------------------------------------- pipeline
    def predict(self, image_path: str, conf_threshold: float = 0.5, nms_threshold: float = 0.4,
                class_names: list[str] | None = None) -> list[Detection]:
        image = self.load_image(image_path)
        batch = self.preprocess_image(image)
        with torch.no_grad():
            out = self.model(batch)
        return self.predict_batch(out[:1], conf_threshold, nms_threshold, class_names)[0]

    def predict_images(self, image_paths, conf_threshold: float = 0.5, nms_threshold: float = 0.4,
                       class_names: list[str] | None = None, batch_size: int = 64, lazy: bool = False) -> list[list[Detection]]:
        """predict() for many images (the loop of predict.py:128-185): every image is preprocessed to the
        same (3, 448, 448) tensor, `batch_size` of them go through the model in one forward, and decode + NMS
        of the whole batch is one launch.  Per image the same result as predict(path)."""
        out: list[list[Detection]] = []
        paths = list(image_paths)
        for lo in range(0, len(paths), max(1, int(batch_size))):
            batch = torch.cat([self.preprocess_image(self.load_image(p)) for p in paths[lo:lo + batch_size]])
            with torch.no_grad():
                pred = self.model(batch)
            out.extend(self.predict_batch(pred, conf_threshold, nms_threshold, class_names, lazy=lazy))
        return out

    def predict_batch(self, predictions: torch.Tensor, conf_threshold: float = 0.5, nms_threshold: float = 0.4,
                      class_names: list[str] | None = None, lazy: bool = False) -> list[list[Detection]]:
        """Decode + NMS for a whole (N,S,S,B*5+C) batch in one launch; one list per image.

        Same result per image as parse_predictions() followed by non_max_suppression() -- including
        the error: the reference builds a validated Detection for EVERY candidate above the confidence
        threshold before NMS runs (inference.py:133-136, 197-208), so a candidate outside [0, 1] raises
        pydantic's ValidationError even when NMS would have suppressed it.  The kernel flags such
        images (`bad_range`); the first flagged image is then decoded without NMS, which builds its
        Detections in the reference's cell / box order and raises at the same candidate.

        lazy=True (extension): each image's result is a LazyDetections -- a read-only sequence over the kernel's
        arrays that builds a validated `Detection` only when an element is first read (pydantic construction, ~10 us
        per object, is the host bottleneck of a large batch).  Compares equal to the eager list."""
        S, B = self.model.S, self.model.B
        Cn = predictions.shape[-1] - B * 5
        predictions = self._on_cuda(predictions)
        r = head.decode_nms(predictions, conf_threshold, nms_threshold, head.IOU_INFER,
                            head.ORDER_GLOBAL_CONF, S, B, Cn)
        bad = r.bad_range.cpu()
        if bool(bad.any()):
            first = int(torch.nonzero(bad)[0].item())
            one = predictions[first] if predictions.dim() == 4 else predictions
            self.parse_predictions(one, conf_threshold, class_names)   # raises ValidationError
            raise _lib.YoloHeadError("yh_decode_nms flagged an out-of-range candidate that the schema accepted")
        return self._build_lazy(r, class_names) if lazy else self._build(r, class_names)

    def parse_predictions(self, pred: torch.Tensor, conf_threshold: float,
                          class_names: list[str] | None = None) -> list[Detection]:
        """(S,S,B*5+C) tensor -> every candidate above the threshold, in cell/box order
        (inference.py:141-210).  Raises pydantic's ValidationError when a kept candidate
        is outside [0,1], as the reference's schema does."""
        S, B = self.model.S, self.model.B
        Cn = pred.shape[-1] - B * 5
        r = head.decode_nms(self._on_cuda(pred), conf_threshold, float("nan"), head.IOU_INFER,
                            head.ORDER_GLOBAL_CONF, S, B, Cn)
        dets = self._build(r, class_names, sort_by_candidate=True)
        return dets[0]

    def iou(self, bbox1: BoundingBox, bbox2: BoundingBox) -> float:
        """IoU with +1e-6 in the denominator (inference.py:212-249)."""
        a = (bbox1.x, bbox1.y, bbox1.width, bbox1.height)
        b = (bbox2.x, bbox2.y, bbox2.width, bbox2.height)
        return float(head.iou_pairs([a], [b], head.IOU_INFER, self._cuda_device()).item())

    def non_max_suppression(self, detections: list[Detection], nms_threshold: float = None,
                            iou_threshold: float = None) -> list[Detection]:
        """Greedy same-class NMS in confidence order (inference.py:251-317)."""
        if iou_threshold is not None:
            warnings.warn("Parameter 'iou_threshold' is deprecated, use 'nms_threshold' instead.",
                          DeprecationWarning, stacklevel=2)
            threshold = iou_threshold
        elif nms_threshold is not None:
            threshold = nms_threshold
        else:
            threshold = 0.4
        if len(detections) == 0:
            return []
        n_keep, keep_idx = head.nms_lists(
            [0, len(detections)], [d.class_id for d in detections], [d.confidence for d in detections],
            [(d.bbox.x, d.bbox.y, d.bbox.width, d.bbox.height) for d in detections], threshold,
            head.IOU_INFER, head.ORDER_GLOBAL_CONF, self._cuda_device())
        n = int(n_keep[0].item())
        return [detections[i] for i in keep_idx[:n].cpu().tolist()]

    # ---------------------------------------------------------------- internals
    def _cuda_device(self) -> torch.device:
        d = torch.device(self.device)
        if d.type != "cuda":
            if not torch.cuda.is_available():
                raise _lib.YoloHeadError("no CUDA device: YOLOInference post-processing has no CPU fallback")
            d = torch.device("cuda", torch.cuda.current_device())
        return d

    def _on_cuda(self, t: torch.Tensor) -> torch.Tensor:
        return t if t.is_cuda else t.to(self._cuda_device())

    @staticmethod
    def _build_lazy(r: head.DecodeResult, class_names) -> "list[LazyDetections]":
        """One LazyDetections per image over host copies of the kernel's outputs (four device->host copies per batch)."""
        n_keep = r.n_keep.cpu().numpy()
        cls, conf, box = r.keep_cls.cpu().numpy(), r.keep_conf.cpu().numpy(), r.keep_box.cpu().numpy()
        return [LazyDetections(cls[i, :n], conf[i, :n], box[i, :n], class_names) for i, n in enumerate(n_keep.tolist())]

    @staticmethod
    def _build(r: head.DecodeResult, class_names, sort_by_candidate: bool = False) -> list[list[Detection]]:
        n_keep = r.n_keep.cpu().tolist()
        k, cls = r.keep_k.cpu(), r.keep_cls.cpu()
        conf, box = r.keep_conf.cpu(), r.keep_box.cpu()
        out = []
        for i, n in enumerate(n_keep):
            ks, cs = k[i, :n].tolist(), cls[i, :n].tolist()
            fs, bs = conf[i, :n].tolist(), box[i, :n].tolist()
            order = sorted(range(n), key=ks.__getitem__) if sort_by_candidate else range(n)
            dets = []
            for q in order:
                cid = int(cs[q])
                name = class_names[cid] if class_names else f"class_{cid}"
                dets.append(Detection(class_id=cid, class_name=name, confidence=fs[q],
                                      bbox=BoundingBox(x=bs[q][0], y=bs[q][1], width=bs[q][2], height=bs[q][3])))
            out.append(dets)
        return out
