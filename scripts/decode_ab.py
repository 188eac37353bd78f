"""decode timings (graph replay, L2 flushed by memset + read) for A/B runs of library variants:
B = 32 / 64 at 80 x 128 x 128, B = 8 at 80 x 256 x 256, and B = 32 with class-concentrated scores (spill list)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from detectron2_centernet_b200 import decode as D
dev = "cuda:0"
C, K = 80, 100
fl_w = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
fl_r = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
g = torch.Generator(device=dev).manual_seed(1)
cases = []
for tag, B, H in (("B32", 32, 128), ("B64", 64, 128), ("1024^2 B8", 8, 256)):
    cases.append((tag, B, H, torch.randn((B, C, H, H), device=dev, generator=g) * 1.5 - 2.19))
B, H = 32, 128
conc = 0.7 * torch.randn((B, C, H, H), device=dev, generator=g) + (torch.randn((B, C, 1, 1), device=dev, generator=g) * 0.6 - 2.19)
cases.append(("B32 class-concentrated", B, H, conc))
for tag, B, H, hm in cases:
    wh = torch.rand((B, 2, H, H), device=dev, generator=g) * 64
    reg = torch.rand((B, 2, H, H), device=dev, generator=g)
    call = lambda: D.ctdet_decode_batch(hm, wh, reg, down_ratio=4, K=K, logits=True)
    for _ in range(3):
        out = call()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream()
    with torch.cuda.stream(cap):
        call(); cap.synchronize()
        with torch.cuda.graph(graph, stream=cap):
            call()
    torch.cuda.synchronize()
    ts = []
    for _ in range(80):
        fl_w.zero_(); fl_r.sum()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); graph.replay(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    t = float(np.median(ts))
    chk = float(out[1].double().sum()) + float(out[3].double().sum())
    t = float(np.mean(ts))                 # (single event pairs are quantised to ~2 us on this box: the mean resolves finer)
    print("%-24s mean %.2f us (median %.2f, min %.2f)  %.0f GB/s  checksum %.6f" % (tag, t, float(np.median(ts)), min(ts), B * C * H * H * 4 / t / 1e3, chk), flush=True)
