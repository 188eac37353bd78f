"""Times BASELINE configs 3 and 4 (parity cases, not bench lines): decode at B = 64, and target-gen +
decode at 1024^2 (B = 8, 128 objects per image).  CUDA events, L2 flushed between iterations, calls replayed from a CUDA graph."""
import sys, json, numpy as np, torch
sys.path.insert(0, '.')
from detectron2_centernet_b200 import decode as D, target as T, synthetic as S
dev = 'cuda:0'
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timed(call, reps=20):
    # replayed from a CUDA graph like bench.py's step (the kernels are ~100 us, the Python wrapper is not free)
    for _ in range(3): call()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream()
    with torch.cuda.stream(cap):
        call()
        with torch.cuda.graph(graph, stream=cap):
            call()
    torch.cuda.synchronize()
    fn = graph.replay
    for _ in range(3): fn()
    tot = 0.0
    for _ in range(reps):
        flush.zero_(); torch.cuda.synchronize()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / reps * 1e3                      # us
g = torch.Generator(device=dev).manual_seed(3)
out = {}
for name, (B, H) in (('cfg3_decode_B64_512', (64, 128)), ('cfg4_decode_B8_1024', (8, 256))):
    heat = torch.sigmoid(torch.randn((B, 80, H, H), device=dev, generator=g) * 1.5 - 2.19).clamp_(1e-4, 1 - 1e-4)
    wh = torch.rand((B, 2, H, H), device=dev, generator=g) * 64
    reg = torch.rand((B, 2, H, H), device=dev, generator=g)
    us = timed(lambda: D.ctdet_decode_batch(heat, wh, reg, down_ratio=4, K=100))
    out[name] = {'us': round(us, 1), 'GBps': round(heat.numel() * 4 / us / 1e3, 1), 'images_per_s': round(B / us * 1e6)}
bl, cl = S.make_batch_objects(1238, 8, 1024, n=128)
boxes, classes, offsets = S.pack_objects(bl, cl)
b, c, o = [torch.from_numpy(a).to(dev) for a in (boxes, classes, offsets)]
us = timed(lambda: T.target_gen(b, c, o, 8, 80, 256, 256))
out['cfg4_target_gen_B8_1024'] = {'us': round(us, 1), 'GBps': round(8 * 80 * 256 * 256 * 4 / us / 1e3, 1), 'images_per_s': round(8 / us * 1e6)}
print(json.dumps(out))
