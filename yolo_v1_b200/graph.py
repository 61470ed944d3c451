"""CapturedHeadStep -- one CUDA graph for a whole head step on fixed tensors.

A validation / benchmarking loop that runs the same shapes over and over spends a
visible share of each step (about 0.2 of 1.4 ms at 65,536 images) in Python between
the ~25 kernel launches of

    loss, parts = criterion(predictions, targets);  loss.backward()
    metric.reset();  metric.update(predictions, targets);  results = metric.compute()

Every call on that path only enqueues work on the current stream (include/yolohead.h:
no allocation, no synchronisation, sizes and counts stay on the device), so the whole
sequence -- loss forward + backward, decode + NMS + TP/FP, compaction, the multi-GPU
record exchange over peer memory when the metric is a DistributedMAP in "peer" mode,
sort and AP -- can be captured once and replayed with a single launch.  replay() then
costs one graph launch; the 7 KB device->host copy of the results is its last node.

The tensors are captured by address: write new data INTO `predictions` / `targets`
(`.copy_()`), or let the model write its output there, and call replay().  The
reference has no counterpart (its step is Python loops); results are the same dicts
`YOLOLoss.forward` and `mAPMetric.compute` return.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from .loss import _KEYS, YOLOLoss, _YoloLossFn
from .metrics import mAPMetric


class CapturedHeadStep:
    def __init__(self, criterion: YOLOLoss, metric, predictions: torch.Tensor, targets: torch.Tensor,
                 backward: bool = True, warmup: int = 2, fused: bool | None = None, pipelined: bool = False):
        _lib.require_cuda(predictions, "predictions")
        _lib.require_cuda(targets, "targets")
        self.criterion, self.metric = criterion, metric
        self.predictions, self.targets = predictions, targets
        self.backward = backward
        # fused: the loss rides on the update kernel (yh_map_update_loss: one read of the two tensors instead of two).
        # Default: wherever that kernel exists (S=7, B=2, C=20, fp32 predictions).
        inner = metric if isinstance(metric, mAPMetric) else metric.metric
        can_fuse = ((inner.S, inner.B, inner.num_classes) == (7, 2, 20) and (criterion.S, criterion.B, criterion.C) == (7, 2, 20)
                    and predictions.dtype is torch.float32)
        if fused and not can_fuse:
            raise _lib.YoloHeadError("CapturedHeadStep(fused=True) needs S=7, B=2, C=20 and fp32 predictions")
        self.fused = can_fuse if fused is None else bool(fused)
        self.grad: Optional[torch.Tensor] = None        # d loss / d predictions after replay()
        self._dist = not isinstance(metric, mAPMetric)  # a DistributedMAP wrapping one
        if self._dist and metric.mode not in ("peer", "auto"):
            raise _lib.YoloHeadError("CapturedHeadStep needs the peer-memory exchange (mode 'peer' or 'auto'): the "
                                     "NCCL exchange reads its split sizes on the host")
        self.recaptures = 0
        # pipelined: TWO graphs of the same step (own record / result buffers each) replayed in turn, so that
        # submit() can queue step k + 1 before collect() has decoded step k: the GPU never waits for Python between
        # steps.  Both run on the same private streams, hence strictly one after the other on the device.
        self._slots = 2 if pipelined else 1
        self._queue: list = []                          # submitted, not yet collected slots (oldest first)
        self._next = 0
        self._host: Optional[torch.Tensor] = None       # pinned landing buffer of the packed results (eager runs)
        dev = predictions.device
        # Two branches: the metric chain is captured on a high-priority stream, the loss on a normal
        # one.  The loss kernel's 25,088 short CTAs would otherwise sit in front of every kernel of
        # the metric chain queued after it; with the priorities the chain's CTAs take the slots the
        # loss CTAs free, and the latency-bound sort passes run inside the bandwidth-bound loss.
        self._main = torch.cuda.Stream(device=dev, priority=-1)
        self._side = torch.cuda.Stream(device=dev, priority=0)
        # eager runs first: workspaces, peer blocks and lazily created state exist before the capture
        side = self._main
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            if self._dist:        # one plain compute() resolves "auto" and sizes the receive blocks
                metric.reset()
                metric.update(predictions.detach(), targets)
                metric.compute()
                if metric.mode != "peer":
                    raise _lib.YoloHeadError("peer-memory exchange unavailable on this box: run the step eagerly")
            for _ in range(max(warmup, 1)):
                self._eager()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._capture()

    def _capture(self):
        lib = _lib.lib()
        # every workspace handed out during a capture has its address baked into the graph: hold them
        # (the caches in _lib.Workspace are per stream, and these two streams are private to this object,
        # so nothing else regrows them; the references make that independent of the caches' lifetime)
        self._held: list = []
        self._graphs, self._outs, self._hosts, self._grads, self._done = [], [], [], [], []
        for _ in range(self._slots):
            graph = torch.cuda.CUDAGraph()
            host = torch.empty(self._host.numel(), dtype=torch.float64, pin_memory=True)
            before = lib.yh_launch_count()
            with _lib.Workspace.holding(self._held), torch.cuda.graph(graph, stream=self._main):
                out = self._enqueue()
                host.copy_(out, non_blocking=True)   # the read-back is a node of the graph too
            self.launches_per_replay = int(lib.yh_launch_count() - before)   # this library's kernels in one graph
            self._graphs.append(graph)
            self._outs.append(out)
            self._hosts.append(host)
            self._grads.append(self.grad)
            self._done.append(torch.cuda.Event())
        self.graph = self._graphs[0]
        self._queue, self._next = [], 0

    # ---------------------------------------------------------------- the step, enqueue only
    def _inner(self) -> mAPMetric:
        return self.metric.metric if self._dist else self.metric

    def _enqueue(self) -> torch.Tensor:
        """The step as two branches that join at the end: the metric chain on the current stream and
        the loss (which shares nothing with it but its inputs) on a side stream forked after the
        update kernel.  The loss kernel is bandwidth-bound while compaction, the record exchange
        (flag waits, NVLink writes) and the small sort-setup kernels are latency-bound, so they
        overlap instead of queueing behind each other."""
        c = self.criterion
        dev = self.predictions.device
        main = torch.cuda.current_stream(dev)
        if self.fused:
            # the loss rides on the update kernel; its tail (object cells, fold) is forked to the side stream so that
            # compaction / exchange / sort start right behind the update kernel
            self.metric.reset()
            parts, self.grad = self.metric.update_with_loss(self.predictions.detach(), self.targets, c, self.backward,
                                                            tail_stream=self._side)
            vec = self.metric.enqueue_peer(dev) if self._dist else self.metric._enqueue_reduce()
            main.wait_stream(self._side)                 # join
            return torch.cat([vec, parts.to(torch.float64)])
        p = self.predictions.detach().requires_grad_(self.backward)
        self.metric.reset()
        self.metric.update(p.detach(), self.targets)
        self._side.wait_stream(main)                     # fork (forking before the update kernel measures the same)
        with torch.cuda.stream(self._side):
            total, parts = _YoloLossFn.apply(p, self.targets, c.S, c.B, c.C, c.lambda_coord, c.lambda_noobj)
            if self.backward:
                total.backward()
                self.grad = p.grad
            parts64 = parts.to(torch.float64)
        if self._dist:
            vec = self.metric.enqueue_peer(dev)
        else:
            vec = self.metric._enqueue_reduce()
        main.wait_stream(self._side)                     # join
        return torch.cat([vec, parts64])

    def _decode(self, host: torch.Tensor):
        loss = dict(zip(_KEYS, host[-5:].tolist()))
        if self._dist:
            return loss, self.metric.finish_peer(host[:-5])
        return loss, self.metric._results_from_packed(host[:-5])

    def _eager(self):
        out = self._enqueue()
        if self._host is None or self._host.numel() != out.numel():   # allocated outside the capture
            self._host = torch.empty(out.numel(), dtype=torch.float64, pin_memory=True)
        loss, res = self._decode(out.cpu())
        if res is None:           # receive blocks too small for this data: the eager compute() grows them
            res = self.metric.compute()
        return loss, res

    # ---------------------------------------------------------------- public
    def submit(self) -> None:
        """Queue one step on whatever `predictions` / `targets` hold when the GPU gets to it (stream order: data
        written into them on the current stream before this call is what the step sees).  Returns at once."""
        if len(self._queue) >= self._slots:
            raise _lib.YoloHeadError(f"CapturedHeadStep: {len(self._queue)} step(s) in flight and not collected; "
                                     "call collect() first" + ("" if self._slots > 1 else " (or pipelined=True)"))
        i = self._next
        self._next = (i + 1) % self._slots
        self._graphs[i].replay()
        self._done[i].record(torch.cuda.current_stream(self.predictions.device))
        self._queue.append(i)

    def collect(self) -> Tuple[Dict[str, float], Dict[str, float]]:
        """Wait for the oldest submitted step and return its (loss_dict, metric results), as criterion(...)[1] and
        metric.compute() would.  `grad` is then that step's d loss / d predictions."""
        if not self._queue:
            raise _lib.YoloHeadError("CapturedHeadStep.collect(): nothing was submitted")
        i = self._queue.pop(0)
        self._done[i].synchronize()
        self.grad = self._grads[i]
        loss, res = self._decode(self._hosts[i])
        if res is None:
            # A skewed class mix overflowed the peer receive blocks (every rank sees it, nothing was moved).  Drain
            # what is in flight, redo the step eagerly on the tensors' current contents -- the eager compute() takes
            # the NCCL exchange and re-creates larger blocks (collective) -- and capture again against them.
            while self._queue:
                self._done[self._queue.pop(0)].synchronize()
            dev = self.predictions.device
            self._main.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(self._main):
                loss, res = self._eager()
            torch.cuda.current_stream(dev).wait_stream(self._main)
            torch.cuda.synchronize(dev)
            self._capture()
            self.recaptures += 1
        return loss, res

    def replay(self) -> Tuple[Dict[str, float], Dict[str, float]]:
        """Run the captured step on whatever `predictions` / `targets` hold now and wait for it:
        submit() + collect().  Returns (loss_dict, metric results)."""
        if self._queue:
            raise _lib.YoloHeadError("CapturedHeadStep.replay(): steps submitted earlier are still uncollected")
        self.submit()
        return self.collect()
