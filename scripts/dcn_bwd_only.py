import sys, math, torch
sys.path.insert(0, '.')
from detectron2_centernet_b200 import deform_conv as DC
dev = 'cuda:0'
B, Cin, Cout, S = 32, int(sys.argv[1]) if len(sys.argv) > 1 else 64, int(sys.argv[2]) if len(sys.argv) > 2 else 64, int(sys.argv[3]) if len(sys.argv) > 3 else 128
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn((B, Cin, S, S), device=dev, generator=g)
off = (torch.rand((B, 18, S, S), device=dev, generator=g) - 0.5) * 4
m = torch.sigmoid(torch.randn((B, 9, S, S), device=dev, generator=g))
w = (torch.rand((Cout, Cin, 3, 3), device=dev, generator=g) - 0.5) * 2 / math.sqrt(9 * Cin)
gy = torch.randn((B, Cout, S, S), device=dev, generator=g)
for _ in range(2): r = DC.dcnv2_backward(x, off, m, w, gy, 1, 1, 1)
torch.cuda.synchronize()
print(float(r[0].abs().mean()))
