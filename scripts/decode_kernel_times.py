"""Per-kernel decode times without ncu: CUDA events around each of N back-to-back calls are too coarse, so this
times the whole decode b2b (graph) for a few env settings passed on the command line by the caller."""
import sys, torch, numpy as np
sys.path.insert(0, '.')
from detectron2_centernet_b200 import decode as D
dev = 'cuda:0'
C, K = 80, 100
flush = torch.empty(512 * 1024 * 1024 // 4, device=dev)
out = []
for B in (32, 64):
    H = W = 128
    g = torch.Generator(device=dev).manual_seed(1)
    hm = torch.randn((B, C, H, W), device=dev, generator=g) * 1.5 - 2.19
    wh = torch.rand((B, 2, H, W), device=dev, generator=g) * 64
    reg = torch.rand((B, 2, H, W), device=dev, generator=g)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(2): D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=True)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=True)
    torch.cuda.synchronize()
    ts = []
    for _ in range(40):
        flush.zero_()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(); gr.replay(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    gr.replay(); s.record()
    for _ in range(50): gr.replay()
    e.record(); torch.cuda.synchronize()
    out.append('B=%d wflush %.1f (mean %.1f) b2b %.1f' % (B, np.median(ts) * 1e3, np.mean(ts) * 1e3, s.elapsed_time(e) / 50 * 1e3))
print(' | '.join(out))
