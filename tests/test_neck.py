"""The DLA-34 up-sampling neck + ctdet heads around the DCNv2 contraction (BASELINE config 5,
SURVEY section 8 rows a11 / a12 / a14).
CPU: the mirror in detectron2_centernet_b200/neck.py against the reference's own IDAUp / DLAUp classes
(same DeformConvV2 stand-in in both, so the comparison is exact).
GPU: the neck with the B200 DeformConvV2 (tcgen05 kernels, bf16 operands) against the same neck with the
fp32 torchvision stand-in, same state_dict; tolerance: relative L2 <= 2e-2 per output map."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import neck_oracle as NO                      # noqa: E402
from oracle import ref_loader as RL                       # noqa: E402

CH = (16, 32, 64, 128, 256, 512)


def _features(B, size, device, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn((B, c, size >> i, size >> i), generator=g).to(device) for i, c in enumerate(CH)]


@pytest.mark.skipif(not RL.available(), reason="reference tree not present (GPU box)")
def test_mirror_equals_reference_neck_structure():
    from detectron2_centernet_b200 import neck as N
    ref = NO.load_reference_neck()
    first = 2
    up_ch = list(CH[first:])
    scales = [2 ** i for i in range(len(up_ch))]
    r_up = ref["DLAUp"](first, list(up_ch), scales)
    r_ida = ref["IDAUp"](CH[first], list(CH[first:5]), [2 ** i for i in range(5 - first)])
    mine = N.DLA34Neck(CH, down_ratio=4, last_level=5, dcn=NO.OracleDeformConvV2)
    # same names, same shapes: the reference state_dict loads into the mirror unchanged
    NO.randomize(r_up, 1); NO.randomize(r_ida, 2)
    mine.dla_up.load_state_dict(r_up.state_dict(), strict=True)
    mine.ida_up.load_state_dict(r_ida.state_dict(), strict=True)
    assert sum(1 for m in mine.modules() if isinstance(m, NO.OracleDeformConvV2)) == 16
    feats = _features(1, 64, "cpu")
    for m in (r_up, r_ida, mine):
        m.eval()
    with torch.no_grad():
        x = r_up([f.clone() for f in feats])
        y = [x[i].clone() for i in range(5 - first)]
        r_ida(y, 0, len(y))
        z = mine([f.clone() for f in feats])
    assert len(y) == len(z) == 3
    for a, b in zip(y, z):
        assert torch.equal(a, b)
    assert z[-1].shape == (1, 64, 16, 16)


def test_fill_up_weights_is_bilinear():
    from detectron2_centernet_b200.neck import fill_up_weights
    up = torch.nn.ConvTranspose2d(3, 3, 4, stride=2, padding=1, groups=3, bias=False)
    fill_up_weights(up)
    k = torch.tensor([0.25, 0.75, 0.75, 0.25])
    assert torch.allclose(up.weight[0, 0], torch.outer(k, k)) and torch.equal(up.weight[2, 0], up.weight[0, 0])


def test_head_shapes_and_bias_prior():
    from detectron2_centernet_b200.neck import CtdetHead
    head = CtdetHead(64, num_classes=80)
    out = head(torch.zeros(1, 64, 8, 8))
    assert set(out) == {"hm", "wh", "reg"} and out["hm"].shape == (1, 80, 8, 8) and out["wh"].shape == (1, 2, 8, 8)
    assert float(head.hm[-1].bias.detach()[0]) == pytest.approx(-2.19) and float(head.wh[-1].bias.detach().abs().max()) == 0.0


@pytest.mark.gpu
def test_neck_with_b200_dcn_matches_fp32_neck_eval():
    """Whole neck, eval mode (BatchNorm + ReLU folded into the DCN epilogue), free-running: the three
    fused maps against the fp32 stand-in neck, then heads + decode on top."""
    from detectron2_centernet_b200 import CenterNetHotPath
    from detectron2_centernet_b200 import neck as N
    dev = "cuda"
    ours = N.DLA34Neck(CH).to(dev)
    ref = N.DLA34Neck(CH, dcn=NO.OracleDeformConvV2).to(dev)
    NO.randomize(ref, 3)
    ours.load_state_dict(ref.state_dict(), strict=True)
    head = N.CtdetHead(64, num_classes=80).to(dev)
    feats = _features(2, 128, dev, seed=5)
    ours.eval(); ref.eval()
    with torch.no_grad():
        a = ours([f.clone() for f in feats]); b = ref([f.clone() for f in feats])
        out = head(a[-1])
        dets = CenterNetHotPath(num_classes=80).inference({k: v for k, v in out.items()}, [(128, 128)] * 2)
    assert len(dets) == 2 and len(dets[0]) <= 100
    for x, y in zip(a, b):
        rel = float((x - y).norm() / y.norm())
        assert rel <= 2e-2, rel                            # bf16 operands, 16 layers deep


@pytest.mark.gpu
def test_neck_train_mode_layers_match_fp32_neck():
    """Train mode (batch-statistics BatchNorm, autograd through dcnv2_backward).  A random-weight neck
    with batch-stat BN amplifies any input perturbation 3-4x per layer, forward and backward (the fp32
    stand-in does the same under fp32 noise: tests/tools/neck_layer_check.py, neck_layer_grad_check.py), so
    a free-running comparison measures the network's conditioning, not the kernels.  The well-posed
    check is per layer: each of the 16 DCNs of the neck (offset conv + contraction) is fed the
    stand-in's own input and output gradient of that layer, and its output and all its gradients are
    compared; DCN + BN + ReLU is compared forward the same way."""
    from detectron2_centernet_b200 import DeformConvV2
    from detectron2_centernet_b200 import neck as N
    dev = "cuda"
    ours = N.DLA34Neck(CH).to(dev)
    ref = N.DLA34Neck(CH, dcn=NO.OracleDeformConvV2).to(dev)
    NO.randomize(ref, 3)
    ours.load_state_dict(ref.state_dict(), strict=True)
    ours.train(); ref.train()
    rec = {}

    def fwd_hook(name):
        def h(m, inp, out):
            rec[name] = {"x": inp[0].detach().clone(), "y": out.detach().clone()}
            out.register_hook(lambda g: rec[name].__setitem__("gy", g.detach().clone()))
        return h

    def in_hook(name):
        def h(m, gin, gout):
            rec[name]["gx"] = gin[0].detach().clone()
        return h

    blocks = [n for n, m in ref.named_modules() if isinstance(m, NO.OracleDeformConvV2)]
    assert len(blocks) == 16
    rmods, omods = dict(ref.named_modules()), dict(ours.named_modules())
    for n in blocks:
        rmods[n].register_forward_hook(fwd_hook(n))
        rmods[n + ".conv"].register_forward_hook(fwd_hook(n + ".conv"))
        rmods[n + ".conv"].register_full_backward_hook(in_hook(n + ".conv"))
    fb = [f.clone().requires_grad_(True) for f in _features(2, 128, dev, seed=5)]
    b = ref(fb)
    b[-1].backward(torch.randn_like(b[-1]))
    rgrads = {k: v.grad.clone() for k, v in ref.named_parameters() if v.grad is not None}

    rel = lambda p, q: float((p - q).norm() / q.norm().clamp_min(1e-20))
    for n in blocks:
        assert isinstance(omods[n], DeformConvV2)
        y = omods[n](rec[n]["x"].clone().requires_grad_(True))            # DCN + batch-stat BN + ReLU
        assert rel(y.detach(), rec[n]["y"]) <= 1e-2, (n, rel(y.detach(), rec[n]["y"]))
        m, r = omods[n + ".conv"], rec[n + ".conv"]                          # the DCN alone, forward + backward
        m.zero_grad()
        x = r["x"].clone().requires_grad_(True)
        y = m(x)
        assert rel(y.detach(), r["y"]) <= 1e-2, (n, rel(y.detach(), r["y"]))
        y.backward(r["gy"])
        assert rel(x.grad, r["gx"]) <= 1e-2, (n, "input", rel(x.grad, r["gx"]))
        for pn, p in m.named_parameters():
            want = rgrads[n + ".conv." + pn]
            if pn == "bias":
                # the gradient that leaves a batch-stat BN sums to zero per channel: compare against the
                # size of what was summed, not against the (rounding-noise) result
                scale = float(r["gy"].abs().sum(dim=(0, 2, 3)).max())
                assert float((p.grad - want).abs().max()) <= 1e-4 * scale, (n, pn)
                continue
            assert rel(p.grad, want) <= 1e-2, (n, pn, rel(p.grad, want))

    # the whole chain through autograd: every parameter and every used base feature gets a finite gradient
    fa = [f.clone().requires_grad_(True) for f in _features(2, 128, dev, seed=5)]
    ours.zero_grad()
    a = ours(fa)
    a[-1].backward(torch.randn_like(a[-1]))
    assert fa[0].grad is None and fa[1].grad is None       # strides 1 and 2 do not feed the neck (first_level = 2)
    assert all(f.grad is not None and bool(torch.isfinite(f.grad).all()) for f in fa[2:])
    assert all(p.grad is not None and bool(torch.isfinite(p.grad).all()) for p in ours.parameters() if p.requires_grad)


def test_upsample_add_cpu_and_autograd_use_the_reference_ops():
    """No GPU / gradients wanted: `upsample_add` is literally `up(x) + skip` (dla.py:175-177)."""
    from detectron2_centernet_b200 import neck as N
    from oracle import neck_oracle as NO
    torch.manual_seed(3)
    up = torch.nn.ConvTranspose2d(8, 8, 4, stride=2, padding=1, groups=8, bias=False)
    N.fill_up_weights(up)
    x, skip = torch.randn(2, 8, 5, 7), torch.randn(2, 8, 10, 14)
    with torch.no_grad():
        assert torch.equal(N.upsample_add(up, x, skip), up(x) + skip)
        assert torch.equal(N.upsample_add(up, x, skip), NO.upsample_add_ref(x, up.weight, skip, 2))
    assert N._is_ida_up(up) and not N._is_ida_up(torch.nn.ConvTranspose2d(8, 8, 4, stride=2, padding=1, bias=False))


@pytest.mark.gpu
@pytest.mark.parametrize("f,shape", [(2, (2, 64, 64, 64)), (2, (3, 24, 7, 9)), (4, (2, 64, 16, 16)), (8, (1, 64, 8, 8)),
                                      (2, (1, 5, 3, 3)), (16, (1, 3, 2, 5))])
def test_upsample_add_kernel_matches_reference_ops(f, shape):
    """ctdet_upsample_add against torch's CPU conv_transpose2d + add: learnable (random) weights and the
    bilinear initialisation, with and without a skip tensor, odd sizes (scalar path).  fp32, 4 products per output in
    a different order than cuDNN / the CPU kernel: tolerance 2e-6 of the output scale."""
    from detectron2_centernet_b200 import neck as N
    from oracle import neck_oracle as NO
    B, C, h, w = shape
    g = torch.Generator().manual_seed(100 + f + h)
    x = torch.randn(shape, generator=g)
    skip = torch.randn((B, C, h * f, w * f), generator=g)
    up = torch.nn.ConvTranspose2d(C, C, 2 * f, stride=f, padding=f // 2, groups=C, bias=False)
    for init in ("bilinear", "random"):
        if init == "bilinear":
            N.fill_up_weights(up)
        else:
            up.weight.data = torch.randn(up.weight.shape, generator=g)
        up_d = torch.nn.ConvTranspose2d(C, C, 2 * f, stride=f, padding=f // 2, groups=C, bias=False).to("cuda:0")
        up_d.weight.data.copy_(up.weight.data)
        with torch.no_grad():
            for sk in (skip, None):
                ref = NO.upsample_add_ref(x, up.weight.data, sk, f)
                got = N.upsample_add(up_d, x.to("cuda:0"), None if sk is None else sk.to("cuda:0")).cpu()
                assert got.shape == ref.shape
                tol = 2e-6 * float(ref.abs().max())
                assert float((got - ref).abs().max()) <= tol, (init, sk is None, float((got - ref).abs().max()), tol)
    # with autograd the torch ops run and gradients flow to the up-sampling weights
    xg = x.to("cuda:0").requires_grad_(True)
    y = N.upsample_add(up_d, xg, skip.to("cuda:0"))
    y.sum().backward()
    assert xg.grad is not None and up_d.weight.grad is not None


@pytest.mark.gpu
@pytest.mark.parametrize("f,shape", [(2, (2, 64, 64, 64)), (2, (3, 24, 7, 9)), (4, (2, 64, 16, 16)), (4, (1, 5, 3, 7)),
                                      (2, (1, 5, 3, 3))])
def test_upsample_add_backward_matches_torch_autograd(f, shape):
    """ctdet_upsample_add_backward (the training path of IDAUp's `up_i(x) + layers[i-1]`, dla.py:175-177) against
    torch autograd through conv_transpose2d + add on the CPU in fp64: grad_x, grad_weight (learnable up-sampling
    weights), grad_skip; with and without a skip, odd sizes.  fp32 sums of 4*f*f products (grad_x) and of B*h*w
    products in atomic order (grad_weight): tolerance 1e-5 of each tensor's scale."""
    from detectron2_centernet_b200 import neck as N
    B, C, h, w = shape
    g = torch.Generator().manual_seed(300 + f + h)
    x = torch.randn(shape, generator=g)
    skip = torch.randn((B, C, h * f, w * f), generator=g)
    gy = torch.randn((B, C, h * f, w * f), generator=g)
    wt = torch.randn((C, 1, 2 * f, 2 * f), generator=g)
    for with_skip in (True, False):
        xr, wr, sr = x.double().requires_grad_(True), wt.double().requires_grad_(True), skip.double().requires_grad_(True)
        y = torch.nn.functional.conv_transpose2d(xr, wr, stride=f, padding=f // 2, groups=C)
        if with_skip:
            y = y + sr
        y.backward(gy.double())
        up = torch.nn.ConvTranspose2d(C, C, 2 * f, stride=f, padding=f // 2, groups=C, bias=False).to("cuda:0")
        up.weight.data.copy_(wt)
        xd = x.to("cuda:0").requires_grad_(True)
        sd = skip.to("cuda:0").requires_grad_(True) if with_skip else None
        out = N.upsample_add(up, xd, sd)
        assert out.grad_fn is not None and "UpsampleAdd" in type(out.grad_fn).__name__      # the fused path, not torch's
        assert float((out.detach().cpu().double() - y.detach()).abs().max()) <= 2e-6 * float(y.abs().max())
        out.backward(gy.to("cuda:0"))
        for name, got, want in (("grad_x", xd.grad, xr.grad), ("grad_weight", up.weight.grad, wr.grad)):
            err = float((got.cpu().double() - want).abs().max())
            assert err <= 1e-5 * float(want.abs().max()), (name, with_skip, err)
        if with_skip:
            assert torch.equal(sd.grad.cpu(), gy)
    # only the input gradient wanted (frozen up-sampling weights, the reference's bilinear init)
    up.weight.requires_grad_(False)
    xd = x.to("cuda:0").requires_grad_(True)
    N.upsample_add(up, xd, None).backward(gy.to("cuda:0"))
    assert float((xd.grad.cpu().double() - xr.grad).abs().max()) <= 1e-5 * float(xr.grad.abs().max())
