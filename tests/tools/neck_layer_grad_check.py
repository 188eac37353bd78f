"""Teacher-forced gradients per DeformConvV2 layer of the neck in train mode, at two levels: the DCN
alone (offset conv + contraction: our kernels) and DCN + batch-stat BN + ReLU."""
import sys, torch
sys.path.insert(0, '.')
from oracle import neck_oracle as NO
from detectron2_centernet_b200 import neck as N
dev = 'cuda'
CH = (16, 32, 64, 128, 256, 512)
g = torch.Generator().manual_seed(5)
f = [torch.randn((2, c, 128 >> i, 128 >> i), generator=g).to(dev) for i, c in enumerate(CH)]
ours = N.DLA34Neck(CH).to(dev); ref = N.DLA34Neck(CH, dcn=NO.OracleDeformConvV2).to(dev)
NO.randomize(ref, 3); ours.load_state_dict(ref.state_dict(), strict=True)
ours.train(); ref.train()
rec = {}
def fh(name):
    def h(m, inp, out):
        rec[name] = {'x': inp[0].detach().clone(), 'y': out.detach().clone()}
        out.register_hook(lambda gg: rec[name].__setitem__('gy', gg.detach().clone()))
    return h
def bh(name):
    def h(m, gin, gout): rec[name]['gx'] = gin[0].detach().clone()
    return h
names = []
for n, m in ref.named_modules():
    if isinstance(m, (NO.OracleDeformConvV2, NO.OracleDCN)):
        m.register_forward_hook(fh(n)); m.register_full_backward_hook(bh(n)); names.append(n)
fb = [t.clone().requires_grad_(True) for t in f]
b = ref(fb); b[-1].backward(torch.randn_like(b[-1]))
rg = {k: v.grad.clone() for k, v in ref.named_parameters() if v.grad is not None}
rel = lambda p, q: float((p - q).norm() / q.norm().clamp_min(1e-20))
om = dict(ours.named_modules())
for n in names:
    m = om[n]; m.zero_grad()
    x = rec[n]['x'].clone().requires_grad_(True)
    y = m(x); y.backward(rec[n]['gy'])
    print('%-26s %4dx%-3d fwd %.4f gx %.4f' % (n, x.shape[1], x.shape[2], rel(y.detach(), rec[n]['y']), rel(x.grad, rec[n]['gx'])),
          {pn.replace('conv_offset_mask', 'om'): round(rel(p.grad, rg[n + '.' + pn]), 4) for pn, p in m.named_parameters()})
