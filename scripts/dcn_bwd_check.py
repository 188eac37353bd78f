"""dcnv2_backward against torchvision autograd (fp32, CUDA) over the shapes the DLA-34 neck produces."""
import sys, math, torch
sys.path.insert(0, '.')
from torchvision.ops import deform_conv2d
from detectron2_centernet_b200 import deform_conv as DC
dev = 'cuda:0'
g = torch.Generator(device=dev).manual_seed(0)
shapes = [(2, 512, 256, 4, 4), (2, 256, 256, 8, 8), (2, 256, 128, 8, 8), (2, 128, 128, 16, 16), (2, 128, 64, 16, 16),
          (2, 64, 64, 32, 32), (2, 256, 64, 8, 8), (1, 64, 64, 5, 7), (2, 64, 64, 8, 16), (2, 64, 64, 16, 8), (3, 128, 64, 12, 20)]
for B, Cin, Cout, H, W in shapes:
    x = torch.randn((B, Cin, H, W), device=dev, generator=g, requires_grad=True)
    off = ((torch.rand((B, 18, H, W), device=dev, generator=g) - 0.5) * 4).requires_grad_(True)
    m = torch.sigmoid(torch.randn((B, 9, H, W), device=dev, generator=g)).requires_grad_(True)
    w = ((torch.rand((Cout, Cin, 3, 3), device=dev, generator=g) - 0.5) * 2 / math.sqrt(9 * Cin)).requires_grad_(True)
    b = torch.zeros(Cout, device=dev, requires_grad=True)
    gy = torch.randn((B, Cout, H, W), device=dev, generator=g)
    y = deform_conv2d(x, off, w, b, padding=1, mask=m)
    y.backward(gy)
    ref = [x.grad, off.grad, m.grad, w.grad, b.grad]
    got = DC.dcnv2_backward(x.detach(), off.detach(), m.detach(), w.detach(), gy, 1, 1, 1)
    yo = DC.dcnv2_forward(x.detach(), off.detach(), m.detach(), w.detach(), b.detach(), 1, 1, 1)
    rels = [float((a - r).norm() / r.norm().clamp_min(1e-12)) for a, r in zip(got, ref)]
    print('%s  fwd %.4f  gi %.4f go %.4f gm %.4f gw %.4f gb %.4f' % ((B, Cin, Cout, H, W), float((yo - y).norm() / y.norm()), *rels))
