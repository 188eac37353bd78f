import sys, torch
sys.path.insert(0, '.')
from detectron2_centernet_b200 import decode as D
dev = 'cuda:0'
B, C, H, W, K = int(sys.argv[1]) if len(sys.argv) > 1 else 32, 80, 128, 128, 100
g = torch.Generator(device=dev).manual_seed(1)
hm = torch.randn((B, C, H, W), device=dev, generator=g) * 1.5 - 2.19
wh = torch.rand((B, 2, H, W), device=dev, generator=g) * 64
reg = torch.rand((B, 2, H, W), device=dev, generator=g)
for _ in range(4):
    out = D.ctdet_decode_batch(hm, wh, reg, 4, False, K, logits=True)
torch.cuda.synchronize()
print(out[1][0, :5])
