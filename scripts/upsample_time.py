"""ctdet_upsample_add per shape: every (C, h, w, f) the DLA-34 neck calls it with at 512^2, batch 32 (collected by
running the neck once), timed alone with CUDA events (L2 flushed by a 256 MB memset before every call), with the
algorithmic bytes x + skip + out, next to the reference's two torch ops (cuDNN grouped transposed conv + add)."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from detectron2_centernet_b200 import neck as N
dev = torch.device("cuda", 0)
torch.manual_seed(0)
B = 32
net = N.DLA34Neck().to(dev).eval()
feats = [torch.randn(B, c, 512 >> i, 512 >> i, device=dev) for i, c in enumerate((16, 32, 64, 128, 256, 512))]
calls = []
orig = N.upsample_add
def spy(up, x, skip=None):
    calls.append((up, tuple(x.shape), up.stride[0]))
    return orig(up, x, skip)
N.upsample_add = spy
with torch.no_grad():
    net(feats)
N.upsample_add = orig
flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
def timed(fn, reps=20):
    ts = []
    for _ in range(reps):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return float(np.median(ts)) * 1e3
tot_ours = tot_ref = 0.0
rows = {}
with torch.no_grad():
    for up, shp, f in calls:
        key = (shp[1], shp[2], shp[3], f)
        x = torch.randn(shp, device=dev)
        skip = torch.randn(shp[0], shp[1], shp[2] * f, shp[3] * f, device=dev)
        for _ in range(3): orig(up, x, skip); (up(x) + skip)
        torch.cuda.synchronize()
        if key not in rows:
            nbytes = 4 * (x.numel() + 2 * skip.numel())
            t = timed(lambda: orig(up, x, skip)); r = timed(lambda: up(x) + skip)
            rows[key] = {"C": key[0], "h": key[1], "w": key[2], "f": f, "calls_per_step": 0, "MB": round(nbytes / 1e6, 1),
                         "ours_us": round(t, 1), "ours_GBps": round(nbytes / t / 1e3, 0), "torch_us": round(r, 1)}
        rows[key]["calls_per_step"] += 1
        tot_ours += rows[key]["ours_us"]; tot_ref += rows[key]["torch_us"]
for r in rows.values(): print(json.dumps(r))
print(json.dumps({"per_step_us_ours": round(tot_ours, 1), "per_step_us_torch": round(tot_ref, 1), "calls_per_step": len(calls)}))
