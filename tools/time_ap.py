"""ap_reduce timing only (sort experiments)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import head, synth
n = 65536
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, "uniform", seed=1234, targets=tgt).cuda(); tgt = tgt.cuda()
gt = torch.zeros(80, dtype=torch.int32, device="cuda")
ch = head.map_update(pred, tgt, 0.3, 0.4, [0.5 + 0.05 * i for i in range(10)], gt)
k, f, off = head.compact_records(ch); m = int(off[-1].item())
def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
tc = t(lambda: (k[:m].clone(), f[:m].clone()))
print(os.environ.get("YH_NVCC_EXTRA", "default"), "ap_reduce %.3f ms" % (t(lambda: head.ap_reduce(k[:m].clone(), f[:m].clone(), m, gt, 20, 10)) - tc))
