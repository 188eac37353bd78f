"""Decode timed as a CUDA-graph replay (no Python / launch overhead).
modes: wflush = 512 MB memset before every replay (L2 left full of dirty lines, as bench.py does);
       rflush = 512 MB read before every replay (L2 left full of clean lines);
       b2b    = 20 replays back to back, no flush (inputs of 168+ MB exceed the 126 MB L2)."""
import sys, torch, numpy as np
sys.path.insert(0, '.')
from detectron2_centernet_b200 import decode as D
dev = 'cuda:0'
C, K = 80, 100
flush = torch.empty(512 * 1024 * 1024 // 4, device=dev)
cases = [(32, 128, True), (64, 128, True), (32, 128, False), (8, 256, True), (32, 128, 'tight')]
for B, H, logits in cases:
    W = H
    g = torch.Generator(device=dev).manual_seed(1)
    hm = torch.randn((B, C, H, W), device=dev, generator=g) * (0.1 if logits == 'tight' else 1.5) - 2.19
    if logits == 'tight': logits = True   # a random-init head: logits within +-0.3 of the bias
    if not logits:
        hm = torch.clamp(torch.sigmoid(hm), 1e-4, 1 - 1e-4)
    wh = torch.rand((B, 2, H, W), device=dev, generator=g) * 64
    reg = torch.rand((B, 2, H, W), device=dev, generator=g)
    for _ in range(3): D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=logits)
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        # warm up ON the capture stream: the zero-initialised workspace is cached per stream, and its
        # one-time memset must not be captured into the graph
        D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=logits)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            out = D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=logits)
    torch.cuda.synchronize()
    res = {}
    for mode in ('wflush', 'rflush'):
        ts = []
        for _ in range(30):
            if mode == 'wflush': flush.zero_()
            else: flush.sum()
            s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
            s.record(); gr.replay(); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        res[mode] = np.median(ts) * 1e3
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    gr.replay(); s.record()
    for _ in range(20): gr.replay()
    e.record(); torch.cuda.synchronize()
    res['b2b'] = s.elapsed_time(e) / 20 * 1e3
    t = res['rflush']
    print('B=%d H=%d logits=%d decode wflush %.1f  rflush %.1f  b2b %.1f us | rflush: %.0f GB/s  %.0f img/s' % (
        B, H, logits, res['wflush'], res['rflush'], res['b2b'], B * C * H * W * 4 / t / 1e3, B / t * 1e6))
