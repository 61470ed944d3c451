"""torchrun --nproc-per-node 2 tools/dist_timeline.py : split of DistributedMAP.compute() (synchronised phases)."""
import os, sys, time
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import mAPMetric, synth, head, distributed as D

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
n = 65536
tgt = synth.make_targets(n, seed=1234 + rank, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, "uniform", seed=1234 + rank, targets=tgt).to(dev); tgt = tgt.to(dev)
metric = D.DistributedMAP(mAPMetric(20, conf_threshold=0.3), mode=sys.argv[1] if len(sys.argv) > 1 else "class_shard")
for it in range(6):
    metric.reset(); metric.update(pred, tgt)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = metric.compute()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    if rank == 0 and it >= 3:
        print(f"compute() {1e3 * (t1 - t0):.3f} ms")
def phase(fn):
    t0 = time.perf_counter(); out = fn(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    return out, 1e3 * (t1 - t0), 1e3 * (t2 - t0)
if os.environ.get("YH_KERNEL_TABLE"):
    from torch.profiler import profile, ProfilerActivity
    metric.reset(); metric.update(pred, tgt); dist.barrier(); torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            metric.compute()
        torch.cuda.synchronize()
    if rank == 0:
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=60))
if metric.mode == "peer":
    if rank == 0:
        print("peer capacity", metric._peer.capacity, "fallbacks", metric.peer_fallbacks)
    px = metric._peer
    for it in range(4):
        metric.reset(); metric.update(pred, tgt); dist.barrier(); torch.cuda.synchronize()
        rows = []
        (key, flg, cap, running, gt_count), a, b = phase(lambda: metric._local_records(dev)); rows.append(("compact", a, b))
        (rk, rf, gtg, info), a, b = phase(lambda: px.exchange(key, flg, cap, running, D.class_owners(20, world), gt_count, n)); rows.append(("exchange", a, b))
        (_, _, packed), a, b = phase(lambda: head.ap_reduce(rk, rf, px.capacity, gtg, 20, 10, m_dev=info[:1])); rows.append(("ap_reduce", a, b))
        _, a, b = phase(lambda: px.sum_results(packed, 803, 802)); rows.append(("peer_sum", a, b))
        _, a, b = phase(lambda: torch.cat([packed, info.to(torch.float64)]).cpu()); rows.append(("d2h", a, b))
        if rank == 0 and it == 3:
            print(" | ".join(f"{nm} host {x:.3f} done {y:.3f}" for nm, x, y in rows), "recv", int(info[0]))
    metric.close(); dist.destroy_process_group(); sys.exit(0)
# phase split: host enqueue time / time until the device is idle, per phase
m = metric.metric
for it in range(4):
    metric.reset(); metric.update(pred, tgt); dist.barrier(); torch.cuda.synchronize()
    rows = []
    (key, flg, cap, running, gt_count), a, b = phase(lambda: metric._local_records(dev)); rows.append(("compact", a, b))
    def f_gt():
        g = torch.cat([gt_count.to(torch.int64), torch.tensor([n], dtype=torch.int64, device=dev)]); dist.all_reduce(g); return g
    glob, a, b = phase(f_gt); rows.append(("allreduce_gt", a, b))
    (pk, pf, part), a, b = phase(lambda: D.partition_records(key, flg, cap, D.class_owners(20, world), world, m_dev=running)); rows.append(("partition", a, b))
    (k2, f2), a, b = phase(lambda: D.exchange_by_owner(pk, pf, part)); rows.append(("exchange", a, b))
    gtc = glob[:80].to(torch.int32)
    (_, _, packed), a, b = phase(lambda: head.ap_reduce(k2, f2, k2.numel(), gtc, 20, 10)); rows.append(("ap_reduce", a, b))
    _, a, b = phase(lambda: dist.all_reduce(packed[:803])); rows.append(("allreduce_ap", a, b))
    _, a, b = phase(lambda: packed.cpu()); rows.append(("d2h", a, b))
    if rank == 0 and it == 3:
        print(" | ".join(f"{nm} host {x:.3f} done {y:.3f}" for nm, x, y in rows))
dist.destroy_process_group()
