"""Order check of yh_ap_reduce's sort on raw records (debugging aid): sorts random 50-bit keys and compares the
caller's arrays (where four short-plan passes leave the result) with numpy's stable sort."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_v1_b200 import _lib, head

dev = torch.device("cuda:0")
for m, seed in [(30000, 1), (30000, 2), (300000, 3), (4_000_000, 4)]:
    rng = np.random.default_rng(seed)
    key = np.uint64(0x3FD0_0000_0000_0000) + rng.integers(0, 1 << 50, size=m).astype(np.uint64)
    cls = rng.integers(0, 20, size=m).astype(np.uint64)
    flags = (cls << np.uint64(56)) | rng.integers(0, 1 << 13, size=m).astype(np.uint64)
    g = torch.full((80,), 500, dtype=torch.int32, device=dev)
    k = torch.from_numpy(key.view(np.int64).copy()).to(dev)
    f = torch.from_numpy(flags.view(np.int64).copy()).to(dev)
    head.ap_reduce(k, f, m, g, 20, 3)
    torch.cuda.synchronize()
    ws = head._ap_workspace(dev, _lib.lib().yh_ap_workspace_bytes(m, 20, 3))
    w = ws[:96].cpu().numpy().view(np.uint32)
    order = np.lexsort((np.arange(m), key, cls.astype(np.int64)))
    gk, gf = k.cpu().numpy().view(np.uint64), f.cpu().numpy().view(np.uint64)
    bad = np.nonzero((gk != key[order]) | (gf != flags[order]))[0]
    top = (key >> np.uint64(23)) | (cls << np.uint64(50))
    u, c = np.unique(top, return_counts=True)
    print(f"m={m} err={w[16]} overflow={w[17]} queued={w[22]} colliding_tops={(c > 1).sum()} mismatches={len(bad)}")
    for b in bad[:6]:
        print("   at", b, "got", hex(int(gk[b])), hex(int(gf[b])), "want", hex(int(key[order][b])), hex(int(flags[order][b])))
