"""Whole-decode timeline from %globaltimer stamps (library built with -DCTDET_DECODE_TIMING, e.g. .variants/lib_timing.so
copied over libctdet_b200.so on the GPU box): first / last stamp per event over all CTAs, microseconds from the seed's start."""
import ctypes, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from detectron2_centernet_b200 import decode as D, _lib
lib = _lib.load()
dev = "cuda:0"
C, K = 80, 100
names = ["seed first start", "seed last end", "first bound published", "last bound published", "sweep first CTA start",
         "sweep last CTA start", "first sweep CTA past its bound", "sweep last end", "tail first CTA start",
         "first tail CTA past griddepcontrol.wait", "last tail CTA past the wait", "tail last end"]
fl_w = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
fl_r = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
g = torch.Generator(device=dev).manual_seed(1)
for tag, B, H in (("B=32 512^2", 32, 128), ("B=64 512^2", 64, 128), ("B=8 1024^2", 8, 256)):
    hm = torch.randn((B, C, H, H), device=dev, generator=g) * 1.5 - 2.19
    wh = torch.rand((B, 2, H, H), device=dev, generator=g) * 64
    reg = torch.rand((B, 2, H, H), device=dev, generator=g)
    call = lambda: D.ctdet_decode_batch(hm, wh, reg, down_ratio=4, K=K, logits=True)
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    rows = []
    for _ in range(7):
        fl_w.zero_(); fl_r.sum(); torch.cuda.synchronize()
        lib.ctdet_debug_dec_ts(None, 1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); call(); e.record(); torch.cuda.synchronize()
        out = (ctypes.c_ulonglong * 16)()
        lib.ctdet_debug_dec_ts(out, 0)
        t = np.array(list(out)[:12], dtype=np.float64)
        rows.append(np.concatenate([(t - t[0]) / 1e3, [s.elapsed_time(e) * 1e3]]))
    med = np.median(np.array(rows), axis=0)
    print("== %s   (events: %.1f us)" % (tag, med[-1]))
    for n, v in zip(names, med[:12]):
        print("   %-42s %7.2f us" % (n, v))
