import sys, torch
sys.path.insert(0, '.')
from oracle import neck_oracle as NO
from detectron2_centernet_b200 import neck as N
dev = 'cuda'
CH = (16, 32, 64, 128, 256, 512)
def feats(B, size, seed=5):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn((B, c, size >> i, size >> i), generator=g).to(dev) for i, c in enumerate(CH)]
for mode in ('eval', 'train'):
    for size in (128, 256):
        ours = N.DLA34Neck(CH).to(dev); ref = N.DLA34Neck(CH, dcn=NO.OracleDeformConvV2).to(dev)
        NO.randomize(ref, 3); ours.load_state_dict(ref.state_dict(), strict=True)
        (ours.eval(), ref.eval()) if mode == 'eval' else (ours.train(), ref.train())
        f = feats(2, size)
        fa = [t.clone().requires_grad_(True) for t in f]; fb = [t.clone().requires_grad_(True) for t in f]
        a = ours(fa); b = ref(fb)
        ga = torch.randn_like(a[-1])
        a[-1].backward(ga); b[-1].backward(ga)
        print(mode, size, 'fwd', ['%.4f' % float((x - y).norm() / y.norm()) for x, y in zip(a, b)],
              'grad feats', ['%.3f' % float((p.grad - q.grad).norm() / q.grad.norm()) for p, q in zip(fa[2:], fb[2:])])
        wa = dict(ours.named_parameters()); wb = dict(ref.named_parameters())
        worst = sorted(((float((wa[k].grad - wb[k].grad).norm() / wb[k].grad.norm().clamp_min(1e-20)), k) for k in wa if wa[k].grad is not None and wb[k].grad is not None), reverse=True)[:4]
        print('   worst param grads', [('%.3f' % r, k) for r, k in worst])
