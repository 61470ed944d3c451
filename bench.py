"""bench.py -- YOLOv1 head hot path: images/s on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the whole head over one synthetic batch of 65,536 images
(BASELINE.json configs[1], S=7 B=2 C=20) through the package's public API:
    YOLOLoss forward + backward      (fused loss kernel)
    mAPMetric.reset/update/compute   (decode + per-class NMS + TP/FP + AP reduce)
`value` times it with the batch resident in HBM; `e2e` repeats it with the batch in
pinned host memory (H2D inside the timed region, results read back).  Inputs (770 MB)
exceed the 126 MB L2, so no explicit flush is needed between iterations.

`--impl reference` times the CPU oracle port of the same path (oracle/, the reference
itself is pure Python and cannot travel to the GPU box) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

S, B, C, D = 7, 2, 20, 30
IMG_BYTES = S * S * D * 4  # 5,880
METRIC = "head images/sec (loss fwd+bwd; decode+NMS+mAP)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--images", type=int, default=65536, help="images per GPU per step")
    ap.add_argument("--profile", default="uniform", choices=["uniform", "realistic", "correlated", "ties"])
    ap.add_argument("--conf", type=float, default=0.3)
    ap.add_argument("--nms", type=float, default=0.4)
    ap.add_argument("--cpu-sample", type=int, default=4096, help="images in the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=8, help="mini-batches the e2e step streams from the host")
    ap.add_argument("--no-graph", action="store_true", help="time the eager step only (no CapturedHeadStep)")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the BASELINE.json configs[2] / [4] records")
    ap.add_argument("--no-head-linear", action="store_true", help="skip the fused head Linear + loss record (SURVEY 8(f) row 4)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "class_shard", "gather"],
                    help="N>1: how compute() makes the per-class order global (yolo_v1_b200/distributed.py)")
    return ap.parse_args()


def make_batch(n, profile, seed):
    from yolo_v1_b200 import synth

    tgt = synth.make_targets(n, seed=seed, p_obj=0.055, size_mix=True)
    pred = synth.make_preds(n, profile, seed=seed, targets=tgt)
    return pred, tgt


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock / throttle reasons with NVML while the timed region runs."""

    BAD = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown"}
    NOTE = {0x4: "sw_power_cap"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in {**self.BAD, **self.NOTE}.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.002)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------ CPU baseline (oracle port)
def cpu_pass(pred, tgt, conf, nms, threads):
    """One pass of the head on the host: torch-CPU loss fwd+bwd + C oracle update/compute."""
    from oracle import c_oracle, loss_oracle

    loss_oracle.yolo_loss_and_grad(pred, tgt)
    st = c_oracle.update_batch(pred.numpy(), tgt.numpy(), conf, nms, n_threads=threads)
    c_oracle.compute(st, [0.5 + 0.05 * i for i in range(10)], n_threads=threads)


def cpu_baseline(args, repeats=3):
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    n = args.cpu_sample
    pred, tgt = make_batch(n, args.profile, 1234)
    cpu_pass(pred[:256], tgt[:256], args.conf, args.nms, threads)  # warm
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        cpu_pass(pred, tgt, args.conf, args.nms, threads)
        best = min(best, time.perf_counter() - t0)
    return {"value": n / best, "unit": "images/s", "cores": threads, "kind": "port",
            "sample": f"{n} images of the same workload ({args.profile}); torch-CPU loss fwd+bwd + "
                      f"oracle/yolo_oracle.c decode/NMS/TP-FP/AP with {threads} OpenMP threads; best of {repeats}"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    n = args.cpu_sample
    pred, tgt = make_batch(n, args.profile, 1234)
    for _ in range(max(args.warmup, 1)):
        cpu_pass(pred, tgt, args.conf, args.nms, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_pass(pred, tgt, args.conf, args.nms, threads)
    dt = (time.perf_counter() - t0) / args.steps
    v = n / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 loss / f64 geometry", "data": "synthetic",
        "config": workload_config(args, n),
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": threads, "kind": "port",
                         "sample": f"each step = {n} images of the workload on the host: torch-CPU loss fwd+bwd + "
                                   f"oracle/yolo_oracle.c with {threads} threads"},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def head_linear_record(n, tgt_d, peaks, time_call):
    """SURVEY 8(f) row 4 beside the headline: the head's last Linear(4096 -> 1470) + view fused with the loss forward +
    backward (yh_head_linear_loss, tcgen05 / TMEM / TMA) on n rows of bf16 activations, against what the reference's
    autocast step runs -- torch's Linear (cuBLAS, out_features = 1470) followed by the loss (here this package's loss
    kernel, itself ~100x the reference's) -- and against the same with the layer padded to 1472 outputs.  Tensor roofline:
    useful flops 2 * n * 4096 * 1470 over the measured cuBLAS bf16 peak (burst: the kernel is timed alone)."""
    import torch.nn.functional as F

    from yolo_v1_b200.head_linear import linear_loss
    from yolo_v1_b200.loss import loss_fwd_bwd

    dev = tgt_d.device
    g = torch.Generator(device=dev).manual_seed(7)
    x = (torch.randn(n, 4096, device=dev, generator=g) * 0.1).to(torch.bfloat16)
    w = (torch.randn(1470, 4096, device=dev, generator=g) * 0.02).to(torch.bfloat16)
    b = torch.rand(1470, device=dev, generator=g) * 0.6
    b16 = b.to(torch.bfloat16)
    w_pad = torch.zeros(1472, 4096, device=dev, dtype=torch.bfloat16)
    w_pad[:1470] = w
    b_pad = torch.zeros(1472, device=dev, dtype=torch.bfloat16)
    b_pad[:1470] = b16

    def two_calls():
        pred = F.linear(x, w, b16)
        return loss_fwd_bwd(pred.view(n, 7, 7, 30), tgt_d, 7, 2, 20, 5.0, 0.5, True)

    def two_calls_padded():
        pred = F.linear(x, w_pad, b_pad)[:, :1470].contiguous()
        return loss_fwd_bwd(pred.view(n, 7, 7, 30), tgt_d, 7, 2, 20, 5.0, 0.5, True)

    parts_f = linear_loss(x, w, b, tgt_d)[0]
    parts_2 = two_calls()[0]
    agree = bool(torch.allclose(parts_f, parts_2, rtol=2e-2))     # bf16 predictions on both sides, a few round the other way
    t_f = time_call(lambda: linear_loss(x, w, b, tgt_d))
    t_fp = time_call(lambda: linear_loss(x, w, b, tgt_d, want_pred=True))
    t_2 = time_call(two_calls)
    t_2p = time_call(two_calls_padded)
    t_mm = time_call(lambda: F.linear(x, w_pad, b_pad))
    # the reference's default (non-AMP) step: fp32 operands through three TF32 passes against torch's fp32 Linear + the loss kernel
    del w_pad, b_pad
    x32, w32 = x.float(), w.float()

    def two_calls_fp32():
        pred = F.linear(x32, w32, b)
        return loss_fwd_bwd(pred.view(n, 7, 7, 30), tgt_d, 7, 2, 20, 5.0, 0.5, True)

    parts_3x = linear_loss(x32, w32, b, tgt_d)[0]
    parts_32 = two_calls_fp32()[0]
    agree32 = bool(torch.allclose(parts_3x, parts_32, rtol=1e-5))
    t_3x = time_call(lambda: linear_loss(x32, w32, b, tgt_d), reps=5)
    t_32 = time_call(two_calls_fp32, reps=3)
    del x32, w32
    flops = 2.0 * n * 4096 * 1470
    peak = float(peaks.get("bf16_tflops", 1590.0))
    tf = flops / (t_f * 1e-3) / 1e12
    return {"what": "Linear(4096 -> 1470) + view + loss fwd + d loss / d prediction in one kernel, bf16 operands, fp32 accumulate",
            "rows": n, "ms": t_f, "ms_with_fp32_predictions_written": t_fp,
            "roofline": {"bound": "tensor", "kernel": "head_linear_loss_kernel<bf16, bf16, 2>", "achieved": tf, "peak": peak,
                         "unit": "TFLOP/s", "frac": tf / peak, "useful_flops_per_launch": flops,
                         "issued_flops_per_launch": 2.0 * n * 4096 * 1568,
                         "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops, cuBLAS 8192^3 burst)" if "bf16_tflops" in peaks
                         else "of fallback 1.59 PFLOP/s"},
            "torch_linear_plus_loss_kernel_ms": t_2, "torch_linear_padded_1472_plus_loss_kernel_ms": t_2p,
            "cublas_linear_padded_1472_alone_ms": t_mm, "speedup_vs_two_calls": t_2 / t_f,
            "speedup_vs_two_calls_padded": t_2p / t_f, "loss_parts_agree_with_two_calls": agree,
            "loss_total": float(parts_f[0]),
            "fp32_operands": {"what": "the reference's default (non-AMP) step: fp32 operands, 3xTF32 (three tensor-core passes, fp32's "
                                      "accuracy) against torch's fp32 Linear (allow_tf32 off) + the loss kernel",
                              "ms": t_3x, "torch_fp32_linear_plus_loss_kernel_ms": t_32, "speedup_vs_two_calls": t_32 / t_3x,
                              "loss_parts_agree_with_two_calls_rtol_1e-5": agree32}}


def workload_config(args, images_per_gpu):
    return {"workload": f"YOLOv1 head: loss fwd+bwd + decode/per-class NMS (conf {args.conf}, IoU {args.nms}) + "
                        f"TP/FP + mAP@50:95 on {images_per_gpu} synthetic images per GPU (S=7,B=2,C=20)",
            "profile": args.profile, "images_per_gpu": images_per_gpu,
            "l2": "inputs (770 MB at 65,536 images) exceed the 126 MB L2; no explicit flush"}


# ------------------------------------------------------------------ BASELINE.json configs[2] and configs[4]
def other_configs(args, world, rank, dev, timed, barrier):
    """configs[2]: mAP@50:95 over a VOC2007-test-sized set (4,952 correlated images, conf 0.01), STRONG scaling:
    the set is split over the ranks (contiguous ranges) and every rank gets the global result.
    configs[4]: the 1M-image sweep = 131,072 images per GPU on 8 GPUs (decode + NMS + mAP, weak scaling);
    run with whatever N this job has and labelled with the total."""
    import torch.distributed as dist

    from yolo_v1_b200 import mAPMetric, synth
    from yolo_v1_b200.distributed import DistributedMAP, shard_range

    out = {}
    # -- configs[2]
    n_voc = 4952
    pred, tgt = synth.make_exact(n_voc, seed=50)           # the inputs of tests/golden/map_big.npz
    lo, hi = shard_range(n_voc, rank, world)
    p_d, t_d = pred[lo:hi].to(dev), tgt[lo:hi].to(dev)
    m = mAPMetric(num_classes=C, conf_threshold=0.01, nms_threshold=0.4)
    if world > 1:
        m = DistributedMAP(m, mode=args.exchange)

    def voc_step():
        m.reset()
        if hi > lo:
            m.update(p_d, t_d)
        return m.compute()

    for _ in range(3):
        res = voc_step()
    ms, res = timed(voc_step, 10)
    inner = m.metric if world > 1 else m
    out["configs[2]"] = {"workload": "mAP@50:95 over 4,952 synthetic images (synth.make_exact, conf 0.01, NMS IoU 0.4, "
                                     "10 thresholds), reset + update + compute per evaluation",
                         "scaling": "strong", "n_gpus": world, "ms_per_evaluation": ms,
                         "images_per_s": n_voc / (ms * 1e-3), "mAP50:95": float(res["mAP50:95"]),
                         "mAP50": float(res["mAP50"]), "sha256": inner.result_digest(),
                         "golden": "tests/golden/map_big.npz exact4952 (the reference's own compute() on these inputs: "
                                   "mAP50:95 0.2713631647915797)"}
    if world > 1:
        m.close()
    # -- configs[4]
    n_big = 131072
    sweep = {}
    for prof in ("uniform", "realistic"):
        t_h = synth.make_targets(n_big, seed=77 + rank, p_obj=0.055, size_mix=True)
        p_h = synth.make_preds(n_big, prof, seed=77 + rank, targets=t_h)
        p_b, t_b = p_h.to(dev), t_h.to(dev)
        del p_h, t_h
        mb = mAPMetric(num_classes=C, conf_threshold=args.conf, nms_threshold=args.nms)
        if world > 1:
            mb = DistributedMAP(mb, mode=args.exchange)

        def big_step():
            mb.reset()
            mb.update(p_b, t_b)
            return mb.compute()

        for _ in range(2):
            big_step()
        ms_b, res_b = timed(big_step, 5)
        sweep[prof] = {"ms_per_sweep": ms_b, "images_per_s": world * n_big / (ms_b * 1e-3),
                       "mAP50:95": float(res_b["mAP50:95"])}
        if world > 1:
            mb.close()
        del p_b, t_b, mb
        torch.cuda.empty_cache()
    out["configs[4]"] = {"workload": f"decode + NMS + TP/FP + mAP@50:95 sweep, 131,072 images per GPU x {world} GPU(s) = "
                                     f"{world * n_big} images (1,048,576 on 8 GPUs), conf {args.conf}, eager calls",
                         "scaling": "weak", "n_gpus": world, "total_images": world * n_big, "profiles": sweep}
    return out


# ------------------------------------------------------------------ our arm
def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch.distributed as dist

    from yolo_v1_b200 import YOLOLoss, _lib, head, mAPMetric

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.images
    pred_h, tgt_h = make_batch(n, args.profile, 1234 + rank)
    pred_h, tgt_h = pred_h.pin_memory(), tgt_h.pin_memory()
    pred_d, tgt_d = pred_h.to(dev), tgt_h.to(dev)
    # defer_loss_dict: the five loss floats are read back when first looked at (end of the step)
    # instead of stalling the host inside forward(); see yolo_v1_b200/loss.py:LossDict
    crit = YOLOLoss(defer_loss_dict=True)
    metric = mAPMetric(num_classes=C, conf_threshold=args.conf, nms_threshold=args.nms)
    if world > 1:
        from yolo_v1_b200.distributed import DistributedMAP

        metric = DistributedMAP(metric, mode=args.exchange)

    def head_pass(p, t):
        """loss fwd+bwd and the metric update of one (mini-)batch already on the device."""
        p = p.requires_grad_(True)
        total, parts = crit(p, t)
        total.backward()
        metric.update(p.detach(), t)
        return parts

    def step(p, t):
        metric.reset()
        parts = head_pass(p, t)
        res = metric.compute()
        return float(parts["total"]), float(res["mAP50:95"])  # both read back inside the step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out

    # ---- device-resident throughput
    for _ in range(max(args.warmup, 3)):
        step(pred_d.detach(), tgt_d)
    l0 = _lib.lib().yh_launch_count()
    with ClockSampler(local) as clk:
        ms, (loss_total, map5095) = timed(lambda: step(pred_d.detach(), tgt_d), args.steps)
    launches = (_lib.lib().yh_launch_count() - l0) // args.steps
    eager_ms, step_launch = ms, "eager (one Python call per stage)"
    graph_serial = None
    # The same step captured once as a CUDA graph (yolo_v1_b200.CapturedHeadStep: same kernels, same
    # results, one launch + one read-back per step).  It is the `value` when it is available -- with
    # N > 1 that needs the peer-memory exchange -- and the eager figure stays in the line beside it.
    if not args.no_graph and (world == 1 or metric.mode == "peer"):
        from yolo_v1_b200 import CapturedHeadStep

        captured, why = None, ""
        try:
            captured = CapturedHeadStep(crit, metric, pred_d, tgt_d, pipelined=True)
        except Exception as e:  # noqa: BLE001 - the eager figure above stands, and says why
            why = f"{type(e).__name__}: {e}"
        ok = torch.tensor([1 if captured is not None else 0], device=dev)
        if world > 1:   # every rank replays, or none does
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 1:
            def graph_step():
                loss, res = captured.replay()
                return float(loss["total"]), float(res["mAP50:95"])

            for _ in range(max(args.warmup, 3)):
                g_out = graph_step()
            # same mAP bits as the eager step; the captured step computes the loss inside the update kernel
            # (CapturedHeadStep(fused=True): one read of the inputs), whose five sums are added in another order
            assert g_out[1] == map5095 and abs(g_out[0] - loss_total) <= 2e-6 * abs(loss_total), (g_out, loss_total, map5095)
            serial_ms, serial_out = timed(graph_step, args.steps)     # submit, wait, decode -- one step at a time

            def pipelined_steps(k):
                """k steps, step i + 1 submitted before step i's results are decoded; EVERY step's loss and mAP are
                read back (one step later than in the strict form)."""
                out = None
                captured.submit()
                for i in range(k):
                    if i + 1 < k:
                        captured.submit()
                    loss, res = captured.collect()
                    out = (float(loss["total"]), float(res["mAP50:95"]))
                return out

            assert pipelined_steps(3) == serial_out
            with ClockSampler(local) as clk:
                total_ms, (loss_total, map5095) = timed(lambda: pipelined_steps(args.steps), 1)
            ms = total_ms / args.steps
            graph_serial = {"ms_per_step": serial_ms, "value": world * n / (serial_ms * 1e-3),
                            "what": "the same graph, strictly one step at a time: replay() = submit, wait, decode"}
            launches = captured.launches_per_replay
            step_launch = ("cuda_graph (CapturedHeadStep" + (", loss fused into the update kernel" if captured.fused else "") +
                           "; two alternating graphs, step k+1 submitted before step k's results are decoded, every "
                           "step's loss and mAP read back)")
        else:
            step_launch += f"; graph capture unavailable ({why or 'on another rank'})"
    value = world * n / (ms * 1e-3)

    # ---- result check that can see a wrong sort / exchange / flag.  On the `uniform` workload no class reaches
    # recall 0.1, so every AP is the t=0 sentinel 1/11 whatever the record order.  The `correlated` profile has
    # true positives: one step of it is run eagerly, through the captured graph and -- with N > 1 -- on rank 0
    # alone over all ranks' images; the sha256 over the full AP table, TP total, record count and ground-truth
    # counts (mAPMetric.result_digest) must be the same everywhere.
    chk_n = min(n, 8192)
    chk_pred_h, chk_tgt_h = make_batch(chk_n, "correlated", 4321 + rank)
    chk_pred, chk_tgt = chk_pred_h.to(dev), chk_tgt_h.to(dev)
    inner = metric.metric if world > 1 else metric
    conf_saved = inner.conf_threshold
    inner.conf_threshold = 0.01
    chk_loss, chk_map = step(chk_pred.detach(), chk_tgt)
    check = {"profile": "correlated", "images_per_gpu": chk_n, "conf": 0.01, "loss_total": chk_loss, "mAP50:95": chk_map,
             "sha256": inner.result_digest(), "agree": {}}
    if step_launch.startswith("cuda_graph"):
        chk_step = None
        try:
            chk_step = CapturedHeadStep(crit, metric, chk_pred, chk_tgt)
        except Exception as e:  # noqa: BLE001
            check["agree"]["graph_vs_eager"] = f"capture failed: {type(e).__name__}"
        if chk_step is not None:
            g_loss, g_res = chk_step.replay()
            check["agree"]["graph_vs_eager"] = bool(inner.result_digest() == check["sha256"] and
                                                    abs(float(g_loss["total"]) - chk_loss) <= 2e-6 * abs(chk_loss))
            del chk_step
    if world > 1:   # rank 0 repeats the step alone on the concatenated images of all ranks
        allp = [torch.empty_like(chk_pred) for _ in range(world)] if rank == 0 else None
        allt = [torch.empty_like(chk_tgt) for _ in range(world)] if rank == 0 else None
        dist.gather(chk_pred, allp, dst=0)
        dist.gather(chk_tgt, allt, dst=0)
        if rank == 0:
            solo = mAPMetric(num_classes=C, conf_threshold=0.01, nms_threshold=args.nms)
            for p_r, t_r in zip(allp, allt):
                solo.update(p_r, t_r)
            solo.compute()
            check["agree"]["sharded_vs_one_gpu"] = bool(solo.result_digest() == check["sha256"])
            del solo, allp, allt
    inner.conf_threshold = conf_saved
    del chk_pred, chk_tgt
    assert all(v is True for v in check["agree"].values()), check

    # ---- end to end: host buffers in, results out.  The batch arrives from pinned host memory
    # in `e2e_chunks` mini-batches on a copy stream; each is consumed (loss fwd+bwd, update) as
    # soon as it lands, so the head overlaps the PCIe transfer of the next mini-batch.
    n_chunks = max(1, min(args.e2e_chunks, n))
    bounds = [(i * n) // n_chunks for i in range(n_chunks + 1)]
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream(dev)

    def e2e_step():
        metric.reset()
        copy_stream.wait_stream(main_stream)
        landed = []
        with torch.cuda.stream(copy_stream):
            for lo, hi in zip(bounds, bounds[1:]):
                p = pred_h[lo:hi].to(dev, non_blocking=True)
                t = tgt_h[lo:hi].to(dev, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
                landed.append((p, t, ev))
        losses = []
        for p, t, ev in landed:
            main_stream.wait_event(ev)
            p.record_stream(main_stream)
            t.record_stream(main_stream)
            losses.append((head_pass(p, t), p.shape[0]))
        res = metric.compute()
        total = sum(float(parts["total"]) * m for parts, m in losses) / n  # batch mean from mini-batch means
        return total, float(res["mAP50:95"])

    e2e_step()
    e2e_ms, (e2e_loss, e2e_map) = timed(e2e_step, max(2, args.steps // 2))

    # The same steps with the input pipeline a loader with pinned memory gives: the copies of step k + 1 are
    # queued on the copy stream before step k's results are read, so the sort / AP tail of a step (which needs
    # all of its records) runs under the next step's transfer.  Every step's inputs still cross PCIe inside
    # the timed region and every step's results are read back; two batches are resident at a time.
    def issue_copies():
        landed = []
        with torch.cuda.stream(copy_stream):
            for lo, hi in zip(bounds, bounds[1:]):
                p = pred_h[lo:hi].to(dev, non_blocking=True)
                t = tgt_h[lo:hi].to(dev, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
                landed.append((p, t, ev))
        return landed

    def e2e_prefetched(k_steps):
        out = None
        nxt = issue_copies()
        for s_i in range(k_steps):
            cur, nxt = nxt, (issue_copies() if s_i + 1 < k_steps else None)
            metric.reset()
            losses = []
            for p, t, ev in cur:
                main_stream.wait_event(ev)
                p.record_stream(main_stream)
                t.record_stream(main_stream)
                losses.append((head_pass(p, t), p.shape[0]))
            res = metric.compute()
            out = (sum(float(parts["total"]) * m for parts, m in losses) / n, float(res["mAP50:95"]))
            del cur
        return out

    e2e_pf = None
    try:
        k_pf = max(3, args.steps // 2)
        e2e_prefetched(2)
        copy_stream.wait_stream(main_stream)
        pf_total_ms, (pf_loss, pf_map) = timed(lambda: e2e_prefetched(k_pf), 1)
        assert abs(pf_loss - e2e_loss) <= 1e-6 * max(1.0, abs(e2e_loss)) and pf_map == e2e_map, "prefetched e2e results differ"
        e2e_pf = {"ms_per_step": pf_total_ms / k_pf, "steps": k_pf}
    except Exception as exc:  # the strict one-step-at-a-time figure below stays the e2e value
        e2e_pf = {"error": f"{type(exc).__name__}: {exc}"[:200]}
    # The PCIe floor of the same transfer: every rank copies its 770 MB AT THE SAME TIME (the host side -- one
    # NUMA node feeding all GPUs -- is what limits e2e at N > 1), max over ranks; rank 0 alone beside it.
    def h2d_only():
        pred_h.to(dev, non_blocking=True)
        tgt_h.to(dev, non_blocking=True)

    h2d_only()
    h2d_concurrent_ms, _ = timed(h2d_only, 3)
    h2d_ms = None
    if world == 1:
        h2d_ms = h2d_concurrent_ms
    else:
        if rank == 0:
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                h2d_only()
            e1.record()
            torch.cuda.synchronize()
            h2d_ms = e0.elapsed_time(e1) / 3
        barrier()
    e2e = {"value": world * n / (e2e_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": 2 * n * IMG_BYTES,
           "d2h_bytes_per_step": n_chunks * 5 * 4 + (4 * C * 10 + 3 + 4 * C) * 8, "ms_per_step": e2e_ms,
           "mini_batches": n_chunks, "h2d_only_ms": h2d_ms, "h2d_concurrent_ms": h2d_concurrent_ms,
           "share_of_concurrent_copy_floor": h2d_concurrent_ms / e2e_ms,
           "result_check": {"loss_total": e2e_loss, "mAP50:95": e2e_map}}
    if e2e_pf is not None and "ms_per_step" in e2e_pf:
        # secondary figure; `value` above stays the strict one (a step's copies start after the previous step's read)
        e2e_pf["value"] = world * n / (e2e_pf["ms_per_step"] * 1e-3)
        e2e_pf["what"] = ("same steps, the host->device copies of step k+1 queued before step k's results are read "
                          "(input prefetch depth 1, as a pinned-memory loader does)")
    e2e["input_prefetch_variant"] = e2e_pf

    # ---- e2e, targets as the dataset holds them: boxes + class ids on the host, encoded on the GPU
    # (yh_encode_targets, the batched _encode_target of dataset.py:487-532).  Secondary figure: the
    # reference's loader ships encoded (S,S,30) targets, which is what `e2e` above does.
    from yolo_v1_b200 import encode_targets_csr

    obj = tgt_h[..., 4] > 0                                   # (n, S, S) object cells, slot 0
    idx = obj.nonzero()                                       # image-major, then row, then column
    cellv = tgt_h[idx[:, 0], idx[:, 1], idx[:, 2]].to(torch.float64)
    boxes_h = torch.stack([(idx[:, 2] + cellv[:, 0]) / S, (idx[:, 1] + cellv[:, 1]) / S, cellv[:, 2], cellv[:, 3]], 1)
    cls_h = cellv[:, 10:].argmax(1).to(torch.int32)
    off_h = torch.zeros(n + 1, dtype=torch.int64)
    off_h[1:] = obj.reshape(n, -1).sum(1).cumsum(0)
    boxes_h, cls_h, off_h = boxes_h.contiguous().pin_memory(), cls_h.pin_memory(), off_h.pin_memory()
    enc, bad = encode_targets_csr(boxes_h.to(dev), cls_h.to(dev), off_h.to(dev))
    targets_match = bool(torch.equal(enc, tgt_d)) and not bool(bad.any())
    del enc, obj, idx, cellv

    def e2e_boxes_step():
        metric.reset()
        copy_stream.wait_stream(main_stream)
        landed = []
        with torch.cuda.stream(copy_stream):
            for lo, hi in zip(bounds, bounds[1:]):
                b0, b1 = int(off_h[lo]), int(off_h[hi])
                p = pred_h[lo:hi].to(dev, non_blocking=True)
                bx = boxes_h[b0:b1].to(dev, non_blocking=True)
                cl = cls_h[b0:b1].to(dev, non_blocking=True)
                of = off_h[lo:hi + 1].to(dev, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
                landed.append((p, bx, cl, of, b0, ev))
        losses = []
        for p, bx, cl, of, b0, ev in landed:
            main_stream.wait_event(ev)
            for x in (p, bx, cl, of):
                x.record_stream(main_stream)
            t, _ = encode_targets_csr(bx, cl, of - b0)
            losses.append((head_pass(p, t), p.shape[0]))
        res = metric.compute()
        total = sum(float(parts["total"]) * m for parts, m in losses) / n
        return total, float(res["mAP50:95"])

    e2e_boxes_step()
    eb_ms, (eb_loss, eb_map) = timed(e2e_boxes_step, max(2, args.steps // 2))
    box_bytes = n * IMG_BYTES + boxes_h.numel() * 8 + cls_h.numel() * 4 + (n + n_chunks) * 8
    e2e["boxes_variant"] = {
        "what": "same step with targets shipped as boxes + class ids (36 B per object + 8 B per image) and encoded "
                "on the GPU by yh_encode_targets; predictions still come from the host",
        "value": world * n / (eb_ms * 1e-3), "unit": "images/s", "ms_per_step": eb_ms,
        "h2d_bytes_per_step": int(box_bytes), "encoded_targets_identical": targets_match,
        "result_check": {"loss_total": eb_loss, "mAP50:95": eb_map}}

    # ---- per-stage kernels, timed alone with CUDA events on the launching stream
    stages = {}
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    def time_call(fn, reps=10):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    from yolo_v1_b200.loss import loss_fwd_bwd

    t_loss = time_call(lambda: loss_fwd_bwd(pred_d, tgt_d, S, B, C, 5.0, 0.5, True))
    stages["loss_fwd_bwd"] = {"ms": t_loss, "bytes_per_image": 3 * IMG_BYTES,
                              "gbs": 3 * IMG_BYTES * n / (t_loss * 1e-3) / 1e9}
    gt_count = torch.zeros(4 * C, dtype=torch.int32, device=dev)
    thr = [0.5 + 0.05 * i for i in range(10)]
    t_upd = time_call(lambda: head.map_update(pred_d, tgt_d, args.conf, args.nms, thr, gt_count))
    chunk = head.map_update(pred_d, tgt_d, args.conf, args.nms, thr, gt_count.zero_())
    kbar = float(chunk.n_keep.float().mean().item())
    gbar = float(chunk.n_gt.float().mean().item())
    # read pred + tgt + a 128-byte per-image result, + 24 B per kept detection.  (What the kernel
    # writes is a little less: 8 B of counters per image, 16-byte records + 1 candidate byte per kept box.)
    upd_bytes = 2 * IMG_BYTES + 128 + 24 * kbar
    stages["map_update"] = {"ms": t_upd, "bytes_per_image": upd_bytes, "kept_per_image": kbar, "gt_per_image": gbar,
                            "written_bytes_per_image": 8 + 17 * kbar,
                            "gbs": upd_bytes * n / (t_upd * 1e-3) / 1e9}
    # the loss riding on the update kernel (what CapturedHeadStep replays): pred + tgt read once, gradient written
    t_fused = time_call(lambda: head.map_update_loss(pred_d, tgt_d, args.conf, args.nms, thr, gt_count, 5.0, 0.5, True))
    fused_bytes = 3 * IMG_BYTES + 128 + 24 * kbar
    # the same call's two halves, each alone: the update kernel with the loss on it (the step's dominant launch) and its tail
    t_fused_k = time_call(lambda: head.map_update_loss(pred_d, tgt_d, args.conf, args.nms, thr, gt_count, 5.0, 0.5, True, phases=1))
    t_fused_t = time_call(lambda: head.map_update_loss(pred_d, tgt_d, args.conf, args.nms, thr, gt_count, 5.0, 0.5, True, phases=2))
    stages["map_update_loss_kernel"] = {"ms": t_fused_k, "bytes_per_image": fused_bytes, "gbs": fused_bytes * n / (t_fused_k * 1e-3) / 1e9,
                                        "what": "map_update_kernel<1> alone (yh_map_update_loss phases = 1): reads pred + tgt, writes the "
                                                "gradient and the records; the object cells of the gradient are rewritten by the tail"}
    stages["map_update_loss_tail"] = {"ms": t_fused_t, "what": "loss_object_cells + loss_rows_fold (phases = 2); in the captured step "
                                                               "they run on a side stream under compaction / sort"}
    stages["map_update_loss_fused"] = {"ms": t_fused, "bytes_per_image": fused_bytes, "gbs": fused_bytes * n / (t_fused * 1e-3) / 1e9,
                                       "what": "yh_map_update_loss: loss fwd+bwd inside the update kernel + object-cell kernel + fold; "
                                               "replaces loss_fwd_bwd + map_update (their sum: "
                                               f"{stages['loss_fwd_bwd']['ms'] + t_upd:.3f} ms)"}
    t_dec = time_call(lambda: head.decode_nms(pred_d, args.conf, args.nms, head.IOU_INFER, head.ORDER_GLOBAL_CONF,
                                              want_boxes=False))
    stages["decode_nms_infer"] = {"ms": t_dec, "bytes_per_image": IMG_BYTES + 128,
                                  "gbs": (IMG_BYTES + 128) * n / (t_dec * 1e-3) / 1e9}
    t_decm = time_call(lambda: head.decode_nms(pred_d, args.conf, args.nms, head.IOU_METRICS,
                                               head.ORDER_CLASS_GROUPED, want_boxes=False))
    stages["decode_nms_metrics"] = {"ms": t_decm, "bytes_per_image": IMG_BYTES + 128,
                                    "gbs": (IMG_BYTES + 128) * n / (t_decm * 1e-3) / 1e9}
    key, flg, off = head.compact_records(chunk)
    M = int(off[-1].item())
    # radix passes the sort runs for these keys (csrc/ap_reduce.cu key_plan): bits that vary over the input
    # + 5 class bits; more than 32 of them -> the short plan: four passes over the leading 32 bits, then one
    # sweep over the keys (find_runs) that orders the few runs of equal leading bits
    kk = key[:M]
    k_np = kk.cpu().numpy().view(np.uint64)
    v_bits = int(np.bitwise_or.reduce(k_np) & np.bitwise_or.reduce(~k_np)) if M else 0
    nb = (v_bits.bit_length() - ((v_bits & -v_bits).bit_length() - 1)) if v_bits else 0
    short_plan = 32 < nb + 5 <= 64
    passes = 4 if short_plan else (nb + 5 + 7) // 8

    def ap_only():
        k2, f2 = key[:M].clone(), flg[:M].clone()
        head.ap_reduce(k2, f2, M, gt_count, C, 10)

    t_clone = time_call(lambda: (key[:M].clone(), flg[:M].clone()))
    t_ap = time_call(ap_only) - t_clone
    # key-range read 8 + histogram read 16 + passes x (read 16 + write 16) + find_runs read 8 (short plan)
    # + two flag sweeps 8 + 8
    ap_bytes = M * (8 + 16 + 32 * passes + (8 if short_plan else 0) + 16)
    stages["ap_reduce"] = {"ms": t_ap, "records": M, "radix_passes": passes, "key_bits": nb, "short_plan": short_plan,
                           "bytes": ap_bytes,
                           "gbs": ap_bytes / (t_ap * 1e-3) / 1e9}
    t_cmp = time_call(lambda: head.compact_records(chunk))
    stages["compact_records"] = {"ms": t_cmp, "bytes": M * 32 + n * 4, "gbs": (M * 32 + n * 4) / (t_cmp * 1e-3) / 1e9}

    # dominant kernel = the longest single launch of the step `value` times (the sort's passes are ~0.05 ms each): with the
    # captured step the loss rides on the update kernel, otherwise the update kernel or the loss kernel
    fused_step = step_launch.startswith("cuda_graph") and "fused" in step_launch
    if fused_step:
        dom = "map_update_loss_kernel"
    else:
        dom = max(("loss_fwd_bwd", "map_update"), key=lambda k: stages[k]["ms"])
    kernel_name = {"loss_fwd_bwd": "loss_kernel<float, 0>", "map_update": "map_update_kernel<0>",
                   "map_update_loss_kernel": "map_update_kernel<1>"}[dom]
    traffic = None
    try:  # DRAM bytes per launch from the committed ncu --set full capture of the same workload
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tj.get("images") == n and tj.get("profile") == args.profile:
            traffic = tj["kernels"].get(kernel_name)
    except Exception:
        pass
    if fused_step:
        step_kernel_ms = stages["map_update_loss_fused"]["ms"] + stages["compact_records"]["ms"] + stages["ap_reduce"]["ms"]
    else:
        step_kernel_ms = stages["loss_fwd_bwd"]["ms"] + stages["map_update"]["ms"] + stages["compact_records"]["ms"] + \
            stages["ap_reduce"]["ms"]
    roof = {"bound": "hbm", "kernel": kernel_name,
            "achieved": stages[dom]["gbs"], "peak": peak, "unit": "GB/s", "frac": stages[dom]["gbs"] / peak,
            "traffic": traffic, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": stages[dom]["bytes_per_image"] * n,
            "algorithmic_bytes": "SURVEY.md 8(d): 11,888 B per image + 24 B per kept detection (map_update); "
                                 "17,640 B per image (loss fwd+bwd); 6,008 B per image (decode + NMS); the fused loss + "
                                 "update kernel: pred + tgt read once + gradient written + 128 B + 24 B per kept = "
                                 "17,768 B per image + 24 B per kept",
            "launch_includes": ("map_update_kernel<1> alone (stages.map_update_loss_kernel); with its tail kernels the call takes "
                                "stages.map_update_loss_fused.ms") if fused_step else None,
            "launch_ms": stages[dom]["ms"], "share_of_step_kernel_time": stages[dom]["ms"] / step_kernel_ms,
            "all": {k: {"gbs": v.get("gbs"), "frac": (v["gbs"] / peak) if v.get("gbs") else None, "ms": v["ms"]}
                    for k, v in stages.items()}}
    # the two halves BASELINE.json's metric names, each timed alone on this rank (kernels only, inputs in HBM)
    map_ms = stages["map_update"]["ms"] + stages["compact_records"]["ms"] + stages["ap_reduce"]["ms"]
    map_bytes = stages["map_update"]["bytes_per_image"] * n + stages["compact_records"]["bytes"] + stages["ap_reduce"]["bytes"]
    halves = {
        "value_loss": {"what": "loss fwd+bwd", "value": n / (stages["loss_fwd_bwd"]["ms"] * 1e-3), "unit": "images/s per GPU",
                       "ms": stages["loss_fwd_bwd"]["ms"], "frac": stages["loss_fwd_bwd"]["gbs"] / peak},
        "value_map": {"what": "decode + NMS + TP/FP (map_update) + compact + sort / AP (ap_reduce)",
                      "value": n / (map_ms * 1e-3), "unit": "images/s per GPU", "ms": map_ms,
                      "frac": map_bytes / (map_ms * 1e-3) / 1e9 / peak},
    }

    # ---- the other BASELINE.json configs, after the timed region (records, not the headline)
    other = {}
    if not args.no_other_configs:
        try:
            other = other_configs(args, world, rank, dev, timed, barrier)
        except Exception as exc:  # noqa: BLE001 - the headline line must still print
            other = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    # rank 0 alone, after every collective of this run (the other ranks wait at the barrier below)
    head_linear_rec = None
    if not args.no_head_linear and rank == 0 and (S, B, C) == (7, 2, 20):
        try:
            head_linear_rec = head_linear_record(n, tgt_d, peaks, time_call)
        except Exception as exc:  # noqa: BLE001 - the headline line must still print
            head_linear_rec = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    barrier()

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            cpu = cpu_baseline(args)
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 loss / f64 geometry", "data": "synthetic",
            "config": workload_config(args, n), "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clk.summary(), "roofline": roof, "cpu_baseline": cpu, "stages": stages,
            "result_check": {"loss_total": loss_total, "mAP50:95": map5095, "sensitive": check},
            "value_loss": halves["value_loss"], "value_map": halves["value_map"], "other_configs": other,
        }
        line["head_linear_loss"] = head_linear_rec
        line["step_launch"] = step_launch
        line["eager"] = {"value": world * n / (eager_ms * 1e-3), "ms_per_step": eager_ms}
        line["graph_serial"] = graph_serial
        if world > 1:
            line["exchange"] = {"mode": metric.mode, "overflow_fallbacks": metric.peer_fallbacks}
        print(json.dumps(line))
    if world > 1:
        metric.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
