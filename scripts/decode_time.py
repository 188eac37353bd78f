import sys, torch, numpy as np
sys.path.insert(0, '.')
from detectron2_centernet_b200 import decode as D
dev = 'cuda:0'
C, H, W, K = 80, 128, 128, 100
flush = torch.empty(512 * 1024 * 1024 // 4, device=dev)
for B in (32, 64):
    g = torch.Generator(device=dev).manual_seed(1)
    hm = torch.randn((B, C, H, W), device=dev, generator=g) * 1.5 - 2.19
    wh = torch.rand((B, 2, H, W), device=dev, generator=g) * 64
    reg = torch.rand((B, 2, H, W), device=dev, generator=g)
    for _ in range(3): D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=True)
    ts = []
    for _ in range(20):
        flush.zero_()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(); D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=True); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    t = np.median(ts) * 1e3
    print('B=%d decode %.1f us  %.0f GB/s' % (B, t, B * C * H * W * 4 / t / 1e3))
