"""target_gen -> loss back to back (the order of CenterNet.losses), for `ncu --cache-control none`: how much of the
loss kernel's 335 MB of reads the L2 still serves when the sweep walks the tensors from their end."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from detectron2_centernet_b200 import target as T, losses as L, synthetic as S
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
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
for it in range(6):
    flush.zero_()
    T.target_gen(tb, tc, to, B, C, H, W, out=gt)
    L._call(hm, wh, reg, gt["hm"], gt["reg_mask"], gt["ind"], gt["wh"], gt["reg"], alpha, gt["num_pos"], None, 2, (1.0, 0.1, 1.0), True)
torch.cuda.synchronize()
print("ok")
