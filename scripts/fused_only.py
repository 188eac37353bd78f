"""A few eager launches of the fused target+loss path (for ncu captures and quick timing)."""
import sys, torch, numpy as np
sys.path.insert(0, '.')
from detectron2_centernet_b200 import losses as L, synthetic as S
dev = 'cuda:0'
B, C, H, W = 32, 80, 128, 128
bl, cl = S.make_batch_objects(1236, B, 512)
boxes, classes, offsets = S.pack_objects(bl, cl)
tb, tc, to = [torch.from_numpy(x).to(dev) for x in (boxes, classes, offsets)]
g = torch.Generator(device=dev).manual_seed(1)
out = {"hm": torch.randn((B, C, H, W), device=dev, generator=g) * 1.5 - 2.19,
       "wh": torch.rand((B, 2, H, W), device=dev, generator=g) * 64,
       "reg": torch.rand((B, 2, H, W), device=dev, generator=g)}
flush = torch.empty(512 * 1024 * 1024 // 4, device=dev)
def run():
    return L.ctdet_losses_from_objects(out, tb, tc, to, [1.0], (1.0, 0.1, 1.0))
for _ in range(3): run()
ts = []
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 10):
    flush.zero_()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record(); run(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
print('fused target+loss (2 launches, eager) %.1f us' % (np.median(ts) * 1e3))
