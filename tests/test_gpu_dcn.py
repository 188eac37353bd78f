"""GPU: the tcgen05 DCNv2 forward (through the C ABI) against the C oracle.

Stated tolerance (operands rounded to bf16, fp32 accumulation, K = 9*Cin up to 4608):
  max |got - ref| <= 2e-2 * max |ref|  and relative L2 error <= 1e-2 with the packed-bf16 blend,
  relative L2 error <= 6e-3 with DCN_FLAG_FP32_BLEND."""
import math

import numpy as np
import pytest
import torch

from oracle import dcn_oracle as DO

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(seed, B, Cin, Cout, H, W, s=1, p=1, d=1, off_range=2.0):
    rng = np.random.default_rng(seed)
    Ho = (H + 2 * p - (2 * d + 1)) // s + 1
    Wo = (W + 2 * p - (2 * d + 1)) // s + 1
    x = rng.normal(size=(B, Cin, H, W)).astype(np.float32)
    off = rng.uniform(-off_range, off_range, (B, 18, Ho, Wo)).astype(np.float32)
    m = (1 / (1 + np.exp(-rng.normal(size=(B, 9, Ho, Wo))))).astype(np.float32)
    bound = 1 / math.sqrt(9 * Cin)
    w = rng.uniform(-bound, bound, (Cout, Cin, 3, 3)).astype(np.float32)
    b = rng.normal(size=(Cout,)).astype(np.float32) * 0.1
    return x, off, m, w, b


def _run(x, off, m, w, b, s=1, p=1, d=1, **kw):
    from detectron2_centernet_b200 import deform_conv as DC
    t = [None if a is None else torch.from_numpy(a).to(DEV) for a in (x, off, m, w, b)]
    out = DC.dcnv2_forward(*t, stride=s, padding=p, dilation=d, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _errs(got, ref):
    return np.abs(got - ref).max() / np.abs(ref).max(), np.linalg.norm(got - ref) / np.linalg.norm(ref)


SHAPES = [
    (1, 64, 64, 16, 24, 1, 1, 1, 2.0),      # 384 pixels = 3 full tiles
    (2, 128, 64, 20, 20, 1, 1, 1, 2.0),     # 400 pixels: ragged last tile, 2 K blocks per tap
    (1, 64, 256, 12, 12, 1, 1, 1, 8.0),     # N = 256, offsets far outside the image
    (1, 256, 128, 9, 7, 1, 1, 1, 1.0),      # fewer pixels than one tile
    (1, 64, 64, 17, 15, 2, 1, 1, 2.0),      # stride 2
    (1, 64, 48, 10, 11, 1, 2, 2, 2.0),      # dilation 2, Cout = 48 (TMEM allocation rounded to 64)
    (1, 512, 256, 16, 16, 1, 1, 1, 2.0),    # the deepest DLA layer: K = 4608
]


@pytest.mark.parametrize("cfg", SHAPES)
def test_dcn_matches_oracle(cfg):
    B, Cin, Cout, H, W, s, p, d, rng_ = cfg
    x, off, m, w, b = _case(11, B, Cin, Cout, H, W, s, p, d, rng_)
    ref = DO.dcn_forward(x, off, m, w, b, s, p, d, threads=4)
    got = _run(x, off, m, w, b, s, p, d)
    emax, el2 = _errs(got, ref)
    assert emax <= 2e-2 and el2 <= 1e-2, (emax, el2)
    got32 = _run(x, off, m, w, b, s, p, d, fp32_blend=True)
    emax, el2 = _errs(got32, ref)
    assert emax <= 2e-2 and el2 <= 6e-3, (emax, el2)


def test_dcn_zero_offset_is_conv2d_and_no_bias():
    x, off, m, w, b = _case(3, 2, 64, 64, 24, 24)
    got = _run(x, np.zeros_like(off), np.ones_like(m), w, None)
    ref = torch.nn.functional.conv2d(torch.from_numpy(x), torch.from_numpy(w), None, padding=1).numpy()
    emax, el2 = _errs(got, ref)
    assert el2 <= 6e-3, (emax, el2)


def test_dcn_exact_on_bf16_representable_inputs():
    """Integer-valued inputs/weights, integer offsets, unit mask: every product and the fp32
    accumulation are exact, so the kernel must reproduce the oracle bit for bit -- this pins the
    K ordering, the swizzled shared-memory layout and the UMMA descriptors."""
    rng = np.random.default_rng(5)
    B, Cin, Cout, H, W = 1, 64, 64, 16, 16
    x = rng.integers(-4, 5, (B, Cin, H, W)).astype(np.float32)
    w = rng.integers(-2, 3, (Cout, Cin, 3, 3)).astype(np.float32)
    off = rng.integers(-2, 3, (B, 18, H, W)).astype(np.float32)
    m = np.ones((B, 9, H, W), np.float32)
    b = rng.integers(-3, 4, (Cout,)).astype(np.float32)
    ref = DO.dcn_forward(x, off, m, w, b, 1, 1, 1)
    got = _run(x, off, m, w, b)
    assert np.array_equal(got, ref)


def test_dcn_fused_epilogue_and_module():
    from detectron2_centernet_b200 import deform_conv as DC
    torch.manual_seed(0)
    layer = DC.DeformConvV2(64, 64).to(DEV).eval()
    # give the zero-initialised offset conv and the BatchNorm statistics non-trivial values
    with torch.no_grad():
        layer.conv.conv_offset_mask.weight.normal_(0, 0.05)
        layer.conv.conv_offset_mask.bias.normal_(0, 0.5)
        layer.actf[0].running_mean.normal_(0, 0.2)
        layer.actf[0].running_var.uniform_(0.5, 1.5)
        layer.actf[0].bias.normal_(0, 0.1)
    assert set(k for k, _ in layer.named_parameters()) >= {"conv.weight", "conv.bias", "conv.conv_offset_mask.weight",
                                                           "conv.conv_offset_mask.bias", "actf.0.weight", "actf.0.bias"}
    x = torch.randn(2, 64, 32, 32, device=DEV)
    with torch.no_grad():
        fused = layer(x)                                             # BN + ReLU folded into the epilogue
        offset, mask = layer.conv.offset_mask(x)
    ref = DO.dcn_forward(x.cpu().numpy(), offset.cpu().numpy(), mask.cpu().numpy(),
                         layer.conv.weight.detach().cpu().numpy(), layer.conv.bias.detach().cpu().numpy(), 1, 1, 1, threads=2)
    bn = layer.actf[0]
    scale = (bn.weight / torch.sqrt(bn.running_var + bn.eps)).detach().cpu().numpy()
    shift = (bn.bias - bn.running_mean * bn.weight / torch.sqrt(bn.running_var + bn.eps)).detach().cpu().numpy()
    ref = np.maximum(ref * scale[None, :, None, None] + shift[None, :, None, None], 0)
    emax, el2 = _errs(fused.cpu().numpy(), ref)
    assert el2 <= 1e-2, (emax, el2)
    # the unfused path (autograd on) computes the same thing
    unfused = layer(x.requires_grad_(True)).detach()
    emax, el2 = _errs(unfused.cpu().numpy(), ref)
    assert el2 <= 1e-2, (emax, el2)


def test_dcn_unsupported_and_cpu():
    from detectron2_centernet_b200 import deform_conv as DC
    with pytest.raises(NotImplementedError):
        DC.dcnv2_forward(torch.zeros(1, 64, 8, 8), torch.zeros(1, 18, 8, 8), torch.zeros(1, 9, 8, 8),
                         torch.zeros(64, 64, 3, 3), None, 1, 1)
    with pytest.raises(NotImplementedError):                           # Cin not a multiple of 64
        DC.dcnv2_forward(torch.zeros(1, 32, 8, 8, device=DEV), torch.zeros(1, 18, 8, 8, device=DEV),
                         torch.zeros(1, 9, 8, 8, device=DEV), torch.zeros(64, 32, 3, 3, device=DEV), None, 1, 1)


@pytest.mark.gpu
def test_raw27_entry_equals_the_split_offset_mask_call():
    """dcnv2_forward_raw27 (raw conv_offset_mask output + its bias, sigmoid inside the kernel) against dcnv2_forward fed
    with what DCN.offset_mask builds in torch (chunk / cat / sigmoid, resnet.py:308-311): same kernel, same
    operands, only the sigmoid's last ulp may differ -> outputs agree to 1e-5 of the output scale."""
    from detectron2_centernet_b200 import deform_conv as DC
    torch.manual_seed(5)
    dev = "cuda:0"
    B, Cin, Cout, H, W = 2, 64, 128, 24, 40
    x = torch.randn(B, Cin, H, W, device=dev)
    raw = torch.randn(B, 27, H, W, device=dev) * 1.5
    ob = torch.randn(27, device=dev) * 0.5
    w = torch.randn(Cout, Cin, 3, 3, device=dev) / (9 * Cin) ** 0.5
    b = torch.randn(Cout, device=dev)
    sc, sh = torch.rand(Cout, device=dev) + 0.5, torch.randn(Cout, device=dev)
    full = raw + ob.view(1, 27, 1, 1)
    o1, o2, m = torch.chunk(full, 3, dim=1)
    ref = DC.dcnv2_forward(x, torch.cat((o1, o2), dim=1), torch.sigmoid(m), w, b, 1, 1, 1, scale=sc, shift=sh, relu=True)
    got = DC.dcnv2_forward_raw27(x, raw, w, b, ob, 1, 1, 1, scale=sc, shift=sh, relu=True)
    assert float((got - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
    got2 = DC.dcnv2_forward_raw27(x, full, w, b, None, 1, 1, 1, scale=sc, shift=sh, relu=True)
    assert float((got2 - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
    with pytest.raises(RuntimeError):
        DC.dcnv2_forward_raw27(x, raw[:, :18], w, b)


@pytest.mark.gpu
def test_raw27_entry_matches_the_oracle_directly():
    """dcnv2_forward_raw27 against the C restatement of the reference (oracle/dcn_oracle.c) fed with what the reference
    computes from the raw conv_offset_mask output in fp32 on the CPU (resnet.py:308-311): offset = channels 0..17 + bias,
    mask = sigmoid(channels 18..26 + bias).  Same stated tolerance as the plain entry (bf16 operands, fp32 accumulation):
    max error <= 2e-2 of the output scale, relative L2 <= 1e-2.  Includes offsets that leave the image."""
    from detectron2_centernet_b200 import deform_conv as DC
    rng = np.random.default_rng(21)
    B, Cin, Cout, H, W = 2, 64, 64, 20, 28
    x = rng.normal(size=(B, Cin, H, W)).astype(np.float32)
    raw = (rng.normal(size=(B, 27, H, W)) * 2.0).astype(np.float32)      # offsets up to ~ +-6 px: some taps leave the map
    ob = rng.normal(size=(27,)).astype(np.float32)
    w = (rng.uniform(-1, 1, (Cout, Cin, 3, 3)) / math.sqrt(9 * Cin)).astype(np.float32)
    b = rng.normal(size=(Cout,)).astype(np.float32)
    full = raw + ob.reshape(1, 27, 1, 1)
    offset = np.ascontiguousarray(full[:, :18])
    mask = (1.0 / (1.0 + np.exp(-full[:, 18:].astype(np.float64)))).astype(np.float32)
    ref = DO.dcn_forward(x, offset, mask, w, b, 1, 1, 1)
    dev = "cuda:0"
    t = lambda a: torch.from_numpy(a).to(dev)
    for om_bias, src in ((ob, raw), (None, full)):
        got = DC.dcnv2_forward_raw27(t(x), t(np.ascontiguousarray(src)), t(w), t(b),
                                     None if om_bias is None else t(om_bias), 1, 1, 1).cpu().numpy()
        emax = np.abs(got - ref).max() / np.abs(ref).max()
        el2 = np.linalg.norm(got - ref) / np.linalg.norm(ref)
        assert emax <= 2e-2 and el2 <= 1e-2, (om_bias is None, emax, el2)


def test_packed_weight_path_is_bit_equal_and_the_module_cache_tracks_the_weight_version():
    """`dcnv2_pack_weight` + DCN_FLAG_PACKED_WEIGHT (no counterpart in the reference: its GEMM reads the fp32 weights,
    deform_conv_cuda.cu:904) must give the bits of the per-call pack; the eval-mode module keeps the packed image per
    weight version."""
    from detectron2_centernet_b200 import deform_conv as DC
    g = torch.Generator(device="cpu").manual_seed(5)
    B, Cin, Cout, H, W = 2, 128, 64, 20, 24
    x = torch.randn(B, Cin, H, W, generator=g).to(DEV)
    off = ((torch.rand(B, 18, H, W, generator=g) - 0.5) * 4).to(DEV)
    m = torch.sigmoid(torch.randn(B, 9, H, W, generator=g)).to(DEV)
    w = ((torch.rand(Cout, Cin, 3, 3, generator=g) - 0.5) * 0.1).to(DEV)
    b = torch.randn(Cout, generator=g).to(DEV)
    want = DC.dcnv2_forward(x, off, m, w, b, 1, 1, 1)
    pw = DC.dcnv2_pack_weight(w)
    assert pw.numel() == DC._lib.load().dcnv2_packed_weight_bytes(Cin, Cout)
    assert torch.equal(DC.dcnv2_forward(x, off, m, w, b, 1, 1, 1, packed_weight=pw), want)
    raw = torch.randn(B, 27, H, W, generator=g).to(DEV)
    assert torch.equal(DC.dcnv2_forward_raw27(x, raw, w, b, None, 1, 1, 1, packed_weight=pw),
                       DC.dcnv2_forward_raw27(x, raw, w, b, None, 1, 1, 1))
    with pytest.raises(RuntimeError):
        DC.dcnv2_forward(x, off, m, w, b, 1, 1, 1, packed_weight=pw[:-256])
    # module: the cache follows in-place updates of the weight
    mod = DC.DeformConvV2(Cin, Cout).to(DEV).eval()
    with torch.no_grad():
        mod.conv.conv_offset_mask.weight.normal_(0, 0.05)
        y1 = mod(x)
        k1 = mod._packed_key
        assert torch.equal(mod(x), y1) and mod._packed_key == k1
        mod.conv.weight.mul_(0.5)
        y2 = mod(x)
        assert mod._packed_key != k1
        mod.invalidate_packed_weight()
        assert torch.equal(mod(x), y2)
    assert float((y1 - y2).abs().max()) > 0


@pytest.mark.parametrize("cfg", [(2, 64, 64, 24, 20), (1, 128, 64, 16, 16), (1, 256, 128, 9, 7), (2, 64, 256, 33, 17), (1, 512, 256, 16, 16)])
def test_fused_offset_conv_matches_conv2d_plus_the_raw27_call(cfg):
    """`dcnv2_forward_fused_offsets`: conv_offset_mask (deform_conv.py:505-513, convention resnet.py:308-313) on the tensor
    cores inside the DCN call.  (i) its raw 27 channels against F.conv2d -- exact to fp32 summation order on
    bf16-representable operands (<= 1e-5 of the scale), <= 1e-2 relative L2 on general fp32 operands (bf16 rounding over
    K = 9 Cin); (ii) the DCN output bit-equal to `dcnv2_forward_raw27` fed with those same 27 channels (same main kernel, same
    packed input); (iii) against the unfused pair on cuDNN's fp32 offsets within the forward's stated tolerance, with small
    conv_offset_mask weights as in a trained / freshly initialised DCN (offsets of a few pixels)."""
    from detectron2_centernet_b200 import deform_conv as DC
    B, Cin, Cout, H, W = cfg
    g = torch.Generator(device="cpu").manual_seed(5 + Cin)
    x = torch.randn(B, Cin, H, W, generator=g).to(DEV)
    w = ((torch.rand(Cout, Cin, 3, 3, generator=g) - 0.5) * 2 / math.sqrt(9 * Cin)).to(DEV)
    b = (torch.randn(Cout, generator=g) * 0.1).to(DEV)
    wo = (torch.randn(27, Cin, 3, 3, generator=g) * (1.5 / math.sqrt(9 * Cin))).to(DEV)
    bo = (torch.randn(27, generator=g) * 0.2).to(DEV)
    torch.backends.cudnn.allow_tf32 = False
    try:
        # (i) exactness on bf16-representable operands
        xb, wob = x.bfloat16().float(), wo.bfloat16().float()
        _, raw = DC.dcnv2_forward_fused_offsets(xb, w, b, wob, bo, return_offset_mask=True)
        ref = torch.nn.functional.conv2d(xb.double(), wob.double(), None, 1, 1).float()
        assert float((raw - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
        # general operands
        out, raw = DC.dcnv2_forward_fused_offsets(x, w, b, wo, bo, return_offset_mask=True)
        ref = torch.nn.functional.conv2d(x, wo, None, 1, 1)
        assert float((raw - ref).norm() / ref.norm()) <= 1e-2
        # (ii) same main kernel on the same 27 channels
        want = DC.dcnv2_forward_raw27(x, raw, w, b, bo, 1, 1, 1)
        assert torch.equal(out, want)
        # (iii) against the unfused pair (fp32 offsets from cuDNN)
        unf = DC.dcnv2_forward_raw27(x, ref, w, b, bo, 1, 1, 1)
        emax, el2 = _errs(out.cpu().numpy(), unf.cpu().numpy())
        assert emax <= 2e-2 and el2 <= 1e-2, (cfg, emax, el2)
    finally:
        torch.backends.cudnn.allow_tf32 = True


def test_deformconvv2_eval_uses_the_fused_offset_conv():
    from detectron2_centernet_b200 import deform_conv as DC
    torch.manual_seed(3)
    m = DC.DeformConvV2(64, 64).to(DEV).eval()
    torch.nn.init.normal_(m.conv.conv_offset_mask.weight, std=1.0 / math.sqrt(9 * 64))
    torch.nn.init.normal_(m.conv.conv_offset_mask.bias, std=0.2)
    x = torch.randn(2, 64, 24, 40, device=DEV)
    with torch.no_grad():
        a = m(x)
        m.fuse_offset_conv = False
        b = m(x)
    emax, el2 = _errs(a.cpu().numpy(), b.cpu().numpy())
    assert emax <= 2e-2 and el2 <= 1e-2, (emax, el2)
    assert not torch.equal(a, b)                               # (two different offset convolutions really ran)


@pytest.mark.parametrize("cfg", [(2, 64, 256, 24, 20), (1, 64, 80, 16, 16), (1, 128, 200, 9, 7), (2, 64, 16, 33, 40)])
def test_plain_conv3x3_matches_conv2d(cfg):
    """`ctdet_conv3x3` (the heads' first layer, centernet.py:112-121, on the tensor cores): exact to fp32 summation order on
    bf16-representable operands (<= 1e-5 of the scale), <= 1e-2 relative L2 on general fp32 operands; output channels
    beyond 128 come from a second chunk, counts that are not multiples of 16 from a zero-padded weight tile."""
    from detectron2_centernet_b200 import deform_conv as DC
    B, Cin, Cout, H, W = cfg
    g = torch.Generator(device="cpu").manual_seed(7 + Cout)
    x = torch.randn(B, Cin, H, W, generator=g).to(DEV)
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) / math.sqrt(9 * Cin)).to(DEV)
    b = torch.randn(Cout, generator=g).to(DEV)
    torch.backends.cudnn.allow_tf32 = False
    try:
        xb, wb = x.bfloat16().float(), w.bfloat16().float()
        ref = torch.relu(torch.nn.functional.conv2d(xb.double(), wb.double(), b.double(), 1, 1)).float()
        got = DC.conv3x3(xb, wb, b, relu=True)
        assert float((got - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
        ref = torch.nn.functional.conv2d(x, w, b, 1, 1)
        got = DC.conv3x3(x, w, b)
        assert float((got - ref).norm() / ref.norm()) <= 1e-2
        got = DC.conv3x3(x, w, None)
        ref = torch.nn.functional.conv2d(x, w, None, 1, 1)
        assert float((got - ref).norm() / ref.norm()) <= 1e-2
    finally:
        torch.backends.cudnn.allow_tf32 = True


def test_ctdet_head_eval_uses_the_tensor_core_first_layer():
    from detectron2_centernet_b200 import neck as N
    torch.manual_seed(11)
    head = N.CtdetHead(64, 80, 256).to(DEV).eval()
    x = torch.randn(2, 64, 24, 40, device=DEV)
    with torch.no_grad():
        a = head(x)
        head.fuse_first_conv = False
        b = head(x)
    for k in ("hm", "wh", "reg"):
        assert float((a[k] - b[k]).norm() / b[k].norm()) <= 1e-2, k
        assert not torch.equal(a[k], b[k])
    head.fuse_first_conv = True
    y = head(x.requires_grad_(True))                       # under autograd the torch modules run
    y["hm"].sum().backward()
    assert x.grad is not None
