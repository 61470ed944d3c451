import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import synth
from yolo_v1_b200.loss import loss_fwd_bwd
n = 65536
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, "uniform", seed=1234, targets=tgt).cuda(); tgt = tgt.cuda()
def t(fn, reps=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ms = t(lambda: loss_fwd_bwd(pred, tgt, 7, 2, 20, 5.0, 0.5, True))
print(os.environ.get("YH_NVCC_EXTRA", "default"), "loss %.4f ms  %.0f GB/s" % (ms, 17640 * n / ms / 1e6))
