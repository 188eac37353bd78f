import sys, torch
sys.path.insert(0, '.')
from oracle import neck_oracle as NO
from detectron2_centernet_b200 import DeformConvV2
dev = 'cuda'
for chi, cho, S in ((512, 256, 4), (256, 256, 8), (64, 64, 32), (64, 64, 64)):
    ref = NO.OracleDeformConvV2(chi, cho).to(dev); NO.randomize(ref, 1)
    ours = DeformConvV2(chi, cho).to(dev); ours.load_state_dict(ref.state_dict(), strict=True)
    g = torch.Generator().manual_seed(2)
    x = torch.randn((2, chi, S, S), generator=g).to(dev)
    for mode in ('eval', 'train'):
        (ours.eval(), ref.eval()) if mode == 'eval' else (ours.train(), ref.train())
        xa = x.clone().requires_grad_(True); xb = x.clone().requires_grad_(True)
        a = ours(xa); b = ref(xb)
        pre_a = ours.conv(xa); pre_b = ref.conv(xb)
        print(chi, cho, S, mode, 'out %.4f' % float((a - b).norm() / b.norm()), 'pre-BN %.4f' % float((pre_a - pre_b).norm() / pre_b.norm()),
              'pre-BN centred %.4f' % float(((pre_a - pre_a.mean((0, 2, 3), keepdim=True)) - (pre_b - pre_b.mean((0, 2, 3), keepdim=True))).norm() / (pre_b - pre_b.mean((0, 2, 3), keepdim=True)).norm()),
              'mean|std of pre_b %.3f %.3f' % (float(pre_b.mean()), float(pre_b.std())))
