"""Image-sharded mAP over the GPUs of one box.

The reference is single-process (its "distributed" is one remote GPU), so this
module has no counterpart file there; it extends `mAPMetric` the way the
north-star prescribes: every rank runs the fused update kernel on a CONTIGUOUS
range of the images, and only `compute()` talks to the other ranks.

TP/FP flags are final per image (csrc/map_update.cu), so what has to be global is
just the per-class confidence-sorted order and the ground-truth counts
(metrics.py:382, 415-436).  Two exchange modes, same results bit for bit:

  "gather"       one all-gather of the compact records (16 B each) in rank order
                 + an all-reduce of the 4*C ground-truth counts; every rank then
                 runs yh_ap_reduce on the full record set (the north-star's form).
  "class_shard"  records are stably partitioned by class owner
                 (yh_partition_records) and exchanged with ONE all_to_all_single;
                 each rank sorts/reduces only its own classes and the 4*C*T AP
                 table is summed with an all-reduce (classes a rank does not own
                 contribute exact zeros).  R x fewer bytes per rank, sort sharded.

  "peer"         class_shard with the partition and the all-to-all fused into one kernel
                 that writes every run straight into its owner's receive buffer over
                 peer-mapped memory (yh_exchange_records, CUDA IPC blocks on NVLink);
                 run lengths, ground-truth counts and the image count travel the same
                 way, and the AP table is summed through the same blocks (yh_peer_sum):
                 compute() has ONE host read (the packed result) and no NCCL call.
                 Ranks must share a box.

Tie order: the reference's stable sort leaves equal confidences in image order.
Ranks own contiguous image ranges and both modes concatenate in rank order, so the
stable radix sort of yh_ap_reduce sees the same sequence as a single process.

The collectives are plain `torch.distributed` calls and work on any backend
(NCCL over NVLink on the box; gloo in the CPU tests of the exchange logic).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, head
from .metrics import mAPMetric


def shard_range(n_images: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous image range [lo, hi) of `rank`: ceil(n/world) images each, the tail ranks may be short or empty."""
    per = -(-n_images // world) if world > 0 else n_images
    lo = min(n_images, rank * per)
    return lo, min(n_images, lo + per)


def class_owners(num_classes: int, world: int) -> List[int]:
    """Owner rank of every class for the class-sharded reduce: contiguous blocks whose sizes differ by
    at most one (20 classes over 8 ranks: 3,2,3,2,3,2,3,2 -- no rank idle, none above ceil(C/world))."""
    return [c * world // num_classes if world <= num_classes else c for c in range(num_classes)]


# ------------------------------------------------------------------ exchange plumbing (backend agnostic)
def _all_counts(n_local: int, device, group) -> List[int]:
    world = dist.get_world_size(group)
    mine = torch.tensor([n_local], dtype=torch.int64, device=device)
    out = torch.empty(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out, mine, group=group)
    return [int(v) for v in out.tolist()]


def gather_records(key: torch.Tensor, flags: torch.Tensor, group=None, count: Optional[torch.Tensor] = None):
    """All-gather variable-length (key, flags) record arrays in rank order.

    `count` (device int64 [1], optional) is the number of valid records when the arrays are
    over-allocated.  Returns (key_all, flags_all, counts).  One count exchange (the only host
    read), one padded all_gather_into_tensor of a (2, max_count) block per rank, pads dropped."""
    world = dist.get_world_size(group)
    dev = key.device
    mine = count.reshape(1) if count is not None else torch.tensor([key.numel()], dtype=torch.int64, device=dev)
    allc = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(allc, mine, group=group)
    counts = [int(v) for v in allc.tolist()]
    m = counts[dist.get_rank(group)]
    cap = max(max(counts), 1)
    send = torch.zeros(2, cap, dtype=torch.int64, device=dev)
    send[0, :m] = key[:m]
    send[1, :m] = flags[:m]
    flat = torch.empty(world * 2 * cap, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(flat, send.view(-1), group=group)
    recv = flat.view(world, 2, cap)
    key_all = torch.cat([recv[r, 0, :counts[r]] for r in range(world)])
    flags_all = torch.cat([recv[r, 1, :counts[r]] for r in range(world)])
    return key_all, flags_all, counts


def exchange_by_owner(key: torch.Tensor, flags: torch.Tensor, part_count, group=None):
    """all_to_all of records already laid out as `world` contiguous runs (run r goes to rank r).

    part_count: the run lengths, a host list or a device int64 tensor (then this does the one
    host read both split-size lists need).  Returns (key_mine, flags_mine): runs received from
    rank 0, 1, ... concatenated in rank order."""
    world = dist.get_world_size(group)
    dev = key.device
    send_n = part_count if torch.is_tensor(part_count) else torch.tensor(part_count, dtype=torch.int64, device=dev)
    recv_n = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv_n, send_n, group=group)
    both = torch.cat([send_n, recv_n]).tolist()
    send_counts, recv_counts = [int(v) for v in both[:world]], [int(v) for v in both[world:]]
    total, m = sum(recv_counts), sum(send_counts)
    out_k = torch.empty(total, dtype=torch.int64, device=dev)
    out_f = torch.empty(total, dtype=torch.int64, device=dev)
    dist.all_to_all_single(out_k, key[:m].contiguous(), recv_counts, send_counts, group=group)
    dist.all_to_all_single(out_f, flags[:m].contiguous(), recv_counts, send_counts, group=group)
    return out_k, out_f


def partition_records(key: torch.Tensor, flags: torch.Tensor, m: int, owners: List[int], n_parts: int,
                      m_dev: Optional[torch.Tensor] = None):
    """yh_partition_records: stable partition by class owner.

    m is the record count, or the capacity when m_dev (device int64 [1]) carries the count.
    Returns (key, flags, part_count) with part_count a device int64 [n_parts] tensor."""
    _lib.require_cuda(key, "records")
    dev = key.device
    lib = _lib.lib()
    out_k = torch.empty(max(m, 1), dtype=torch.int64, device=dev)
    out_f = torch.empty(max(m, 1), dtype=torch.int64, device=dev)
    cnt = torch.zeros(n_parts, dtype=torch.int64, device=dev)
    ws = torch.empty(lib.yh_partition_workspace_bytes(int(m)) + 256, dtype=torch.uint8, device=dev)
    tab = (_lib.C.c_uint8 * len(owners))(*owners)
    with _lib.on_device(dev):
        rc = lib.yh_partition_records(key.data_ptr() if m else None, flags.data_ptr() if m else None, int(m),
                                      _lib.ptr(m_dev), tab, len(owners), int(n_parts), out_k.data_ptr(),
                                      out_f.data_ptr(), cnt.data_ptr(), ws.data_ptr(), ws.numel(),
                                      _lib.stream_ptr(dev))
    _lib.check(rc, "yh_partition_records")
    return out_k, out_f, cnt


# ------------------------------------------------------------------ peer-memory exchange
class _DeviceBlock:
    """Raw device memory as a zero-copy tensor source (`torch.as_tensor` reads the interface)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


class PeerUnavailable(_lib.YoloHeadError):
    """The ranks could not map each other's memory (not one box, or CUDA IPC is blocked)."""


class PeerExchange:
    """One receive block per rank, mapped into every rank of the group (all on one box).

    Construction and close() are collective.  exchange() enqueues yh_exchange_records on
    the current stream and returns views of this rank's receive buffer; it never reads
    anything back, and the epoch of the flag protocol is the library's own call counter on the
    device (epoch argument 0), so an exchange can sit inside a captured CUDA graph."""

    def __init__(self, capacity: int, device: torch.device, group=None, wait_ms: int = 0):
        lib = _lib.lib()
        self.group, self.device = group, device
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.capacity = int(max(capacity, 1024))
        self.nbytes = int(lib.yh_peer_block_bytes(self.capacity))
        self._own = _lib.C.c_void_p()
        self._opened: List[int] = []
        handle = (_lib.C.c_ubyte * 64)()
        problem = None
        with _lib.on_device(device):
            if lib.yh_peer_alloc(self.nbytes, _lib.C.byref(self._own), handle) != 0:
                problem = "yh_peer_alloc: " + lib.yh_last_error().decode("utf-8", "replace")
        blocks = [None] * self.world
        blocks[self.rank] = self._own.value
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, None if problem else bytes(handle), group=group)
            with _lib.on_device(device):
                for r, h in enumerate(handles):
                    if r == self.rank or problem:
                        continue
                    if h is None:
                        problem = f"rank {r} could not allocate its block"
                        break
                    p = _lib.C.c_void_p()
                    buf = (_lib.C.c_ubyte * 64).from_buffer_copy(h)
                    if lib.yh_peer_open(buf, _lib.C.byref(p)) != 0:
                        problem = f"yh_peer_open(rank {r}): " + lib.yh_last_error().decode("utf-8", "replace")
                        break
                    blocks[r] = p.value
                    self._opened.append(p.value)
            # every block is mapped everywhere before anyone writes -- or nobody uses the blocks
            ok = torch.tensor([0 if problem else 1], dtype=torch.int32, device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                self._release(collective=False)
                raise PeerUnavailable(problem or "another rank could not map the peer blocks")
        elif problem:
            raise PeerUnavailable(problem)
        self._blocks = (_lib.C.c_void_p * self.world)(*blocks)
        if wait_ms:
            with _lib.on_device(device):
                _lib.check(lib.yh_peer_set_wait_ms(self._own, int(wait_ms)), "yh_peer_set_wait_ms")
        raw = torch.as_tensor(_DeviceBlock(self._own.value, self.nbytes), device=device)
        ko, fo = int(lib.yh_peer_key_offset(self.capacity)), int(lib.yh_peer_flags_offset(self.capacity))
        self.recv_key = raw[ko:ko + 8 * self.capacity].view(torch.int64)
        self.recv_flags = raw[fo:fo + 8 * self.capacity].view(torch.int64)
        self._ws = _lib.Workspace()

    def exchange(self, key, flags, m: int, m_dev, owners: List[int], gt_count: torch.Tensor, n_images: int):
        """Returns (recv_key, recv_flags, gt_count_global int32 [4C], info int64 [4]); see yolohead.h."""
        lib, dev = _lib.lib(), self.device
        ws = self._ws.get(dev, int(lib.yh_exchange_workspace_bytes(int(m))))
        gt_glob = torch.empty_like(gt_count)
        info = torch.empty(4, dtype=torch.int64, device=dev)
        tab = (_lib.C.c_uint8 * len(owners))(*owners)
        with _lib.on_device(dev):
            rc = lib.yh_exchange_records(key.data_ptr() if m else None, flags.data_ptr() if m else None, int(m),
                                         _lib.ptr(m_dev), tab, len(owners), self.world, self.rank, self._blocks,
                                         self.capacity, 0, gt_count.data_ptr(), int(n_images),
                                         gt_glob.data_ptr(), info.data_ptr(), ws.data_ptr(),
                                         ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "yh_exchange_records")
        return self.recv_key, self.recv_flags, gt_glob, info

    def exchange_images(self, chunks, owners: List[int], gt_count: torch.Tensor):
        """yh_exchange_images: the same result as compact + exchange(), straight from the fixed-stride records
        of the update batches (`chunks`: head.UpdateChunk list in update order).  world <= 8."""
        lib, dev = _lib.lib(), self.device
        n = len(chunks)
        VP, I = _lib.C.c_void_p, _lib.C.c_int
        nk = (VP * max(n, 1))(*[ch.n_keep.data_ptr() for ch in chunks])
        rk = (VP * max(n, 1))(*[ch.rec_key.data_ptr() for ch in chunks])
        rf = (VP * max(n, 1))(*[ch.rec_flags.data_ptr() for ch in chunks])
        ni = (I * max(n, 1))(*[int(ch.n) for ch in chunks])
        K = int(chunks[0].rec_key.shape[1]) if n else 98
        total = sum(int(ch.n) for ch in chunks)
        ws = self._ws.get(dev, int(lib.yh_exchange_images_workspace_bytes(total, n)))
        gt_glob = torch.empty_like(gt_count)
        info = torch.empty(4, dtype=torch.int64, device=dev)
        tab = (_lib.C.c_uint8 * len(owners))(*owners)
        with _lib.on_device(dev):
            rc = lib.yh_exchange_images(n, nk, rk, rf, ni, K, tab, len(owners), self.world, self.rank, self._blocks,
                                        self.capacity, 0, gt_count.data_ptr(), gt_glob.data_ptr(), info.data_ptr(),
                                        ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "yh_exchange_images")
        return self.recv_key, self.recv_flags, gt_glob, info

    def sum_results(self, vec: torch.Tensor, n: int, err_index: int) -> None:
        """vec[:n] (float64, in place) summed over the ranks through the peer blocks (yh_peer_sum)."""
        with _lib.on_device(self.device):
            rc = _lib.lib().yh_peer_sum(vec.data_ptr(), int(n), int(err_index), self.world, self.rank, self._blocks,
                                        0, _lib.stream_ptr(self.device))
        _lib.check(rc, "yh_peer_sum")

    def close(self):
        """Collective: unmap the peers' blocks, then free the own one once nobody maps it."""
        self._release(collective=True)

    def _release(self, collective: bool):
        if self._own.value is None:
            return
        lib = _lib.lib()
        torch.cuda.synchronize(self.device)
        self.recv_key = self.recv_flags = None
        with _lib.on_device(self.device):
            for p in self._opened:
                lib.yh_peer_close(p)
            self._opened = []
            if collective and self.world > 1:
                dist.barrier(group=self.group)
            lib.yh_peer_free(self._own)
        self._own = _lib.C.c_void_p()


# ------------------------------------------------------------------ the metric
class DistributedMAP:
    """`mAPMetric` whose `compute()` is global over a process group.

    Each rank calls `update()` with ITS images only -- rank r must hold the r-th
    contiguous block of the evaluation set (`shard_range`) for tie order to match
    the single-process reference.  `compute()` is collective and returns the same
    dict on every rank.  Everything else (`reset`, thresholds, attributes) is the
    wrapped metric's.

    Host round trips per compute(): "peer" has one (the packed result); "class_shard" and
    "gather" add one for the exchange's split sizes.  Everything else (compaction, partition,
    exchange, sort, AP) is enqueued.  "peer" maps CUDA IPC blocks between the ranks on first
    use (collective, ranks on one box); call close() (collective) to release them.
    "auto" = "peer" if that mapping works on every rank, else "class_shard"."""

    def __init__(self, metric: mAPMetric, group=None, mode: str = "class_shard", peer_wait_ms: int = 0):
        if mode not in ("gather", "class_shard", "peer", "auto"):
            raise ValueError(f"unknown exchange mode {mode!r}")
        # peer mode: how long a rank's kernels wait for a peer's flag before giving up (0 = the library's 4 s).
        # An eager compute() lines the ranks up first (a one-element all-reduce on the stream), so ordinary skew
        # -- a short tail shard, a model forward still queued -- never comes near it; what is left is a dead peer.
        self.peer_wait_ms = int(peer_wait_ms)
        self.metric = metric
        self.group = group
        self.mode = mode
        self._peer: Optional[PeerExchange] = None
        self.peer_fallbacks = 0            # compute() calls that overflowed the receive blocks
        self.image_exchange = True         # peer mode: yh_exchange_images (False: compact + yh_exchange_records)

    def close(self):
        if self._peer is not None:
            self._peer.close()
            self._peer = None

    def _peer_blocks(self, dev, cap_local: int, need: Optional[int] = None) -> PeerExchange:
        """The mapped receive blocks, created on first use (collective; one host read).

        Capacity: what a rank receives when classes are spread evenly -- the largest local
        capacity x world x (largest share of classes one rank owns) -- or 1.25 x `need`
        after an overflow.  A skewed class mix that does not fit falls back to the NCCL
        exchange for that call and re-creates the blocks larger."""
        if self._peer is not None and need is None:
            return self._peer
        g = self.group
        world = dist.get_world_size(g)
        want = torch.tensor([cap_local if need is None else need], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(want, op=dist.ReduceOp.MAX, group=g)
        top = int(want.item())
        if need is None:
            owners = class_owners(self.metric.num_classes, world)
            share = max(owners.count(r) for r in range(world)) / float(self.metric.num_classes)
            capacity = int(np.ceil(top * min(float(world), world * share)))
        else:
            capacity = int(np.ceil(1.25 * top))
        if self._peer is not None:
            self._peer.close()
        self._peer = PeerExchange(capacity, dev, g, self.peer_wait_ms)
        return self._peer

    def __getattr__(self, name):
        return getattr(self.metric, name)

    def reset(self):
        self.metric.reset()

    def update(self, predictions: torch.Tensor, targets: torch.Tensor):
        self.metric.update(predictions, targets)

    def update_with_loss(self, predictions: torch.Tensor, targets: torch.Tensor, criterion, backward: bool = True,
                         tail_stream=None):
        return self.metric.update_with_loss(predictions, targets, criterion, backward, tail_stream)

    def _local_records(self, dev):
        """(key, flags, capacity, count) of this rank: all batches compacted into one buffer behind
        a device-side running count -- nothing is read back."""
        m = self.metric
        C = m.num_classes
        if m._lists_live or m._pred_lists:
            raise _lib.YoloHeadError("DistributedMAP: the Python list state was touched on this rank; the sharded "
                                     "compute() runs from the device records only")
        cap = sum(ch.rec_key.numel() for ch in m._chunks)
        key = torch.empty(max(cap, 1), dtype=torch.int64, device=dev)
        flg = torch.empty(max(cap, 1), dtype=torch.int64, device=dev)
        running = torch.zeros(1, dtype=torch.int64, device=dev)
        for ch in m._chunks:
            head.compact_records(ch, key, flg, running)
        gt_count = m._gt_count.clone() if m._gt_count is not None else torch.zeros(4 * C, dtype=torch.int32, device=dev)
        return key, flg, cap, running, gt_count

    def enqueue_peer(self, dev) -> torch.Tensor:
        """Peer-mode compute() up to, not including, the host read: exchange, sort + AP, sum over
        the ranks.  Returns the device vector finish_peer() decodes.  Capturable in a CUDA graph.

        Up to 8 ranks the records go from the update batches straight to their owners
        (yh_exchange_images: one pass); beyond that they are compacted first (yh_exchange_records)."""
        m = self.metric
        world = dist.get_world_size(self.group)
        T, C = len(m.iou_thresholds), m.num_classes
        if m._lists_live or m._pred_lists:
            raise _lib.YoloHeadError("DistributedMAP: the Python list state was touched on this rank; the sharded "
                                     "compute() runs from the device records only")
        cap = sum(ch.rec_key.numel() for ch in m._chunks)
        px = self._peer_blocks(dev, cap)
        owners = class_owners(C, world)
        if world <= 8 and self.image_exchange:
            gt_count = m._gt_count if m._gt_count is not None else torch.zeros(4 * C, dtype=torch.int32, device=dev)
            rk, rf, gt_glob, info = px.exchange_images(m._chunks, owners, gt_count)
        else:
            key, flg, cap, running, gt_count = self._local_records(dev)
            n_images = sum(ch.n for ch in m._chunks)
            rk, rf, gt_glob, info = px.exchange(key, flg, cap, running, owners, gt_count, n_images)
        _, _, packed = head.ap_reduce(rk, rf, px.capacity, gt_glob, C, T, m_dev=info[:1])
        if world > 1:
            px.sum_results(packed, 4 * C * T + 3, 4 * C * T + 2)
        return torch.cat([packed, info.to(torch.float64)])

    def finish_peer(self, vec: torch.Tensor) -> Optional[Dict[str, float]]:
        """Host side of a peer-mode compute(): the result dict, or None if a receive block overflowed."""
        m = self.metric
        T, C = len(m.iou_thresholds), m.num_classes
        host = vec.cpu()
        if host[-1] != 0:
            raise _lib.YoloHeadError("peer exchange: a rank of the group did not show up within the wait limit "
                                     f"({self.peer_wait_ms or 4000} ms; DistributedMAP(peer_wait_ms=...))")
        if host[-2] != 0:
            return None
        if int(host[-3]) == 0:
            return {"mAP50:95": 0.0, "mAP50": 0.0, "mAP75": 0.0, "precision": 0.0, "recall": 0.0}
        ap, tp_total, n_pred, gtc = head.unpack_ap(host[:-4], C, T)
        return m._results(ap, tp_total, n_pred, gtc, int(gtc[0].sum()), T, C)

    def compute(self) -> Dict[str, float]:
        m, g = self.metric, self.group
        world = dist.get_world_size(g)
        dev = m._device if m._device is not None else torch.device("cuda", torch.cuda.current_device())
        T, C = len(m.iou_thresholds), m.num_classes
        n_images = len(m._pred_lists) + sum(c.n for c in m._chunks)
        cap = sum(ch.rec_key.numel() for ch in m._chunks)
        if self.mode == "auto":            # "peer" where the ranks can map each other's memory
            try:
                self._peer_blocks(dev, cap)
                self.mode = "peer"
            except PeerUnavailable as e:   # raised on every rank together
                import warnings

                warnings.warn(f"DistributedMAP: peer-memory exchange unavailable ({e}); using the NCCL exchange")
                self.mode = "class_shard"
        if self.mode == "peer":
            if world > 1 and not torch.cuda.is_current_stream_capturing():
                # the flag waits of the exchange are bounded; order the ranks on the stream first, as NCCL would
                dist.all_reduce(torch.zeros(1, dtype=torch.int32, device=dev), group=g)
            res = self.finish_peer(self.enqueue_peer(dev))
            if res is not None:
                return res
            self.peer_fallbacks += 1           # a receive block was too small: NCCL path now, larger blocks next
        key, flg, cap, running, gt_count = self._local_records(dev)
        # ground-truth counts and the image count travel in one all-reduce
        glob = torch.cat([gt_count.to(torch.int64),
                          torch.tensor([n_images], dtype=torch.int64, device=dev)])
        dist.all_reduce(glob, group=g)
        gt_count = glob[:4 * C].to(torch.int32)
        if self.mode == "gather":
            key, flg, counts = gather_records(key, flg, g, count=running)
            _, _, packed = head.ap_reduce(key, flg, sum(counts), gt_count, C, T)
        else:
            pk, pf, part = partition_records(key, flg, cap, class_owners(C, world), world, m_dev=running)
            key, flg = exchange_by_owner(pk, pf, part, g)
            if self.mode == "peer":
                self._peer_blocks(dev, cap, need=key.numel())
            _, _, packed = head.ap_reduce(key, flg, key.numel(), gt_count, C, T)
            # ap table, TP total, record count, error flag: classes a rank does not own are exact zeros
            dist.all_reduce(packed[:4 * C * T + 3], group=g)
        host = torch.cat([packed, glob[4 * C:].to(torch.float64)]).cpu()
        if int(host[-1]) == 0:
            return {"mAP50:95": 0.0, "mAP50": 0.0, "mAP75": 0.0, "precision": 0.0, "recall": 0.0}
        ap, tp_total, n_pred, gtc = head.unpack_ap(host[:-1], C, T)
        return m._results(ap, tp_total, n_pred, gtc, int(gtc[0].sum()), T, C)


def evaluate_sharded(model, images: torch.Tensor, targets: torch.Tensor, batch: int = 256, group=None,
                     mode: str = "class_shard", metric: Optional[mAPMetric] = None, **metric_kw) -> Dict[str, float]:
    """evaluate_model (metrics.py:654-714) over an in-memory set, image-sharded over the group.

    `images` / `targets` are the FULL set on every rank (host or device); each rank runs the
    model only on its contiguous `shard_range`.  model=None treats `images` as predictions."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    lo, hi = shard_range(images.shape[0], rank, world)
    dev = torch.device("cuda", torch.cuda.current_device())
    dm = DistributedMAP(metric or mAPMetric(**({"num_classes": 20} | metric_kw)), group, mode)
    if model is not None:
        model.eval()
    try:
        with torch.no_grad():
            for s in range(lo, hi, batch):
                e = min(hi, s + batch)
                x = images[s:e].to(dev, non_blocking=True)
                t = targets[s:e].to(dev, non_blocking=True)
                dm.update(model(x) if model is not None else x, t)
        return dm.compute()
    finally:
        dm.close()
