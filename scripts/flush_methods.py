"""How the L2 flush between timed iterations changes what a kernel is charged for (B = 32, 80 x 128 x 128).
  memset      : 256 MB memset (bench.py's method) -- leaves up to 126 MB of DIRTY lines that are written back to DRAM
                inside the timed region
  memset+read : the same, then a 256 MB read of another buffer -- L2 full of clean, unrelated lines
  read        : only the read
  none        : back-to-back replays (inputs of 335 / 168 MB are larger than L2)"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from detectron2_centernet_b200 import target as T, losses as L, decode as D, synthetic as S
dev = "cuda:0"
B, C, H, W = 32, 80, 128, 128
bl, cl = S.make_batch_objects(1236, B, 512)
boxes, classes, offsets = S.pack_objects(bl, cl)
tb, tc, to = [torch.from_numpy(x).to(dev) for x in (boxes, classes, offsets)]
g = torch.Generator(device=dev).manual_seed(1)
hm = torch.randn((B, C, H, W), device=dev, generator=g) * 1.5 - 2.19
wh = torch.rand((B, 2, H, W), device=dev, generator=g) * 64
reg = torch.rand((B, 2, H, W), device=dev, generator=g)
gt = T.target_gen(tb, tc, to, B, C, H, W)
alpha = L.alpha_tensor([1.0], C, dev)
fl_w = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
fl_r = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
calls = {
    "loss": lambda: L._call(hm, wh, reg, gt["hm"], gt["reg_mask"], gt["ind"], gt["wh"], gt["reg"], alpha, gt["num_pos"], None, 2, (1.0, 0.1, 1.0), True),
    "target_gen": lambda: T.target_gen(tb, tc, to, B, C, H, W, out=gt),
    "decode": lambda: D.ctdet_decode_batch(hm, wh, reg, down_ratio=4, K=100, logits=True),
}
flushes = {"memset": lambda: fl_w.zero_(), "memset+read": lambda: (fl_w.zero_(), fl_r.sum()), "read": lambda: fl_r.sum(), "none": lambda: None}
for name, call in calls.items():
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream()
    with torch.cuda.stream(cap):
        call(); cap.synchronize()
        with torch.cuda.graph(graph, stream=cap):
            call()
    torch.cuda.synchronize()
    for fname, fl in flushes.items():
        ts = []
        for _ in range(25):
            fl()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); graph.replay(); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e) * 1e3)
        print("%-10s flush=%-12s median %.2f us  min %.2f us" % (name, fname, float(np.median(ts)), min(ts)), flush=True)
