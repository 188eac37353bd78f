"""Per-layer view of the neck comparison: every DeformConvV2 of our neck fed the fp32 stand-in's own
input of that layer (teacher forcing) next to the free-running error of the chain."""
import sys, torch
sys.path.insert(0, '.')
from oracle import neck_oracle as NO
from detectron2_centernet_b200 import neck as N
dev = 'cuda'
CH = (16, 32, 64, 128, 256, 512)
def feats(B, size, seed=5):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn((B, c, size >> i, size >> i), generator=g).to(dev) for i, c in enumerate(CH)]
mode = sys.argv[1] if len(sys.argv) > 1 else 'train'
size = int(sys.argv[2]) if len(sys.argv) > 2 else 128
ours = N.DLA34Neck(CH).to(dev); ref = N.DLA34Neck(CH, dcn=NO.OracleDeformConvV2).to(dev)
NO.randomize(ref, 3); ours.load_state_dict(ref.state_dict(), strict=True)
(ours.eval(), ref.eval()) if mode == 'eval' else (ours.train(), ref.train())
rec_ref, rec_ours = {}, {}
def hook(store, name):
    def h(m, inp, out): store[name] = (inp[0].detach().clone(), out.detach().clone())
    return h
for name, m in ref.named_modules():
    if isinstance(m, NO.OracleDeformConvV2): m.register_forward_hook(hook(rec_ref, name))
from detectron2_centernet_b200 import DeformConvV2
for name, m in ours.named_modules():
    if isinstance(m, DeformConvV2): m.register_forward_hook(hook(rec_ours, name))
f = feats(2, size)
fa = [t.clone().requires_grad_(True) for t in f]; fb = [t.clone().requires_grad_(True) for t in f]
a = ours(fa); b = ref(fb)
mods = dict(ours.named_modules())
rel = lambda x, y: float((x - y).norm() / y.norm().clamp_min(1e-20))
for name, (xin, yout) in rec_ref.items():
    chain_in = rel(rec_ours[name][0], xin); chain_out = rel(rec_ours[name][1], yout)
    m = mods[name]
    hs = [h for h in m._forward_hooks]; m._forward_hooks.clear()
    forced = m(xin.clone().requires_grad_(True)).detach()
    # how much of the stand-in's own output moves when its input is perturbed by fp32 noise of bf16 size
    rm = dict(ref.named_modules())[name]; rm._forward_hooks.clear()
    pert = rm((xin * (1 + 0.002 * torch.randn_like(xin))).requires_grad_(True)).detach()
    print('%-24s in %4d x %3d  chain in %.4f out %.4f | forced %.4f | stand-in under 0.2%% input noise %.4f  frac relu-dead %.2f' % (
        name, xin.shape[1], xin.shape[2], chain_in, chain_out, rel(forced, yout), rel(pert, yout), float((yout == 0).float().mean())))
