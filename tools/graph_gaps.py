"""Timeline of one replay of the captured step: per-kernel start / duration and the idle gaps between kernels."""
import json, os, sys, tempfile
import torch
from torch.profiler import ProfilerActivity, profile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import CapturedHeadStep, YOLOLoss, mAPMetric, synth

n = 65536
tgt = synth.make_targets(n, seed=1234, p_obj=0.055, size_mix=True)
pred = synth.make_preds(n, os.environ.get("PROFILE", "uniform"), seed=1234, targets=tgt).cuda()
tgt = tgt.cuda()
step = CapturedHeadStep(YOLOLoss(), mAPMetric(num_classes=20, conf_threshold=float(os.environ.get("CONF", 0.3))), pred, tgt)
for _ in range(5):
    step.replay()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step.replay()
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), "trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
ev.sort(key=lambda e: e["ts"])
# a replay ends with the device->host copy of the results; take the last complete replay
ends = [i for i, e in enumerate(ev) if "DtoH" in e["name"]]
last = ev[ends[-2] + 1:ends[-1] + 1]
t0 = last[0]["ts"]
end_prev, busy = None, 0.0
for e in last:
    gap = 0.0 if end_prev is None else e["ts"] - end_prev
    print(f"{e['ts'] - t0:9.1f} us  dur {e['dur']:8.1f}  gap {gap:6.1f}  stream {e['args'].get('stream', '?'):>3}  {e['name'][:70]}")
    end_prev = max(end_prev or 0, e["ts"] + e["dur"])
print(f"span {end_prev - t0:.1f} us, events {len(last)}")
