"""GPU: the tcgen05 DCNv2 backward (through the C ABI) against the C oracle's restatement of
modulated_deform_conv_cuda_backward (oracle/dcn_oracle.c, pinned on torchvision autograd by
tests/test_dcn_oracle.py).

Stated tolerance (input, weight, grad_output and the column gradient are rounded to bf16, fp32
accumulation): relative L2 error <= 1.5e-2 per gradient tensor, max-abs error <= 4e-2 of the
tensor's max-abs."""
import math

import numpy as np
import pytest
import torch

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
    gy = rng.normal(size=(B, Cout, Ho, Wo)).astype(np.float32)
    return x, off, m, w, gy


def _reference(x, off, m, w, gy, s, p, d):
    from oracle import dcn_oracle as DO
    return DO.dcn_backward(x, off, m, w, gy, s, p, d)      # (grad_input, grad_offset, grad_mask, grad_weight, grad_bias)


def _errs(got, ref):
    return np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30), np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-30)


SHAPES = [
    (1, 64, 64, 16, 24, 1, 1, 1, 2.0),      # 3 full tiles, one K-block group
    (2, 128, 64, 20, 20, 1, 1, 1, 2.0),     # ragged tiles, 18 K blocks -> 3 groups
    (1, 64, 128, 12, 12, 1, 1, 1, 8.0),     # Cout = 128 (two 64-channel atoms per M tile), offsets far outside
    (1, 128, 256, 9, 7, 1, 1, 1, 1.0),      # Cout = 256: two M tiles; fewer pixels than one tile
    (1, 64, 64, 17, 15, 2, 1, 1, 2.0),      # stride 2
    (1, 64, 64, 10, 11, 1, 2, 2, 2.0),      # dilation 2
]


@pytest.mark.parametrize("cfg", SHAPES)
def test_dcn_backward_weight_and_bias(cfg):
    from detectron2_centernet_b200 import deform_conv as DC
    B, Cin, Cout, H, W, s, p, d, rng_ = cfg
    x, off, m, w, gy = _case(21, B, Cin, Cout, H, W, s, p, d, rng_)
    ref = _reference(x, off, m, w, gy, s, p, d)
    t = [torch.from_numpy(a).to(DEV) for a in (x, off, m, w, gy)]
    _, _, _, gw, gb = DC.dcnv2_backward(t[0], t[1], t[2], t[3], t[4], s, p, d, need_input_grads=False)
    torch.cuda.synchronize()
    emax, el2 = _errs(gw.cpu().numpy(), ref[3])
    assert el2 <= 1.5e-2 and emax <= 4e-2, ("grad_weight", emax, el2)
    emax, el2 = _errs(gb.cpu().numpy(), ref[4])
    assert el2 <= 1e-5, ("grad_bias", emax, el2)


@pytest.mark.parametrize("cfg", SHAPES)
def test_dcn_backward_input_offset_mask(cfg):
    from detectron2_centernet_b200 import deform_conv as DC
    B, Cin, Cout, H, W, s, p, d, rng_ = cfg
    x, off, m, w, gy = _case(22, B, Cin, Cout, H, W, s, p, d, rng_)
    ref = _reference(x, off, m, w, gy, s, p, d)
    t = [torch.from_numpy(a).to(DEV) for a in (x, off, m, w, gy)]
    gi, go, gm, _, _ = DC.dcnv2_backward(t[0], t[1], t[2], t[3], t[4], s, p, d, need_weight_grad=False)
    torch.cuda.synchronize()
    for name, got, r in (("grad_input", gi, ref[0]), ("grad_offset", go, ref[1]), ("grad_mask", gm, ref[2])):
        emax, el2 = _errs(got.cpu().numpy(), r)
        assert el2 <= 1.5e-2 and emax <= 4e-2, (name, emax, el2)


def test_dcn_autograd_module():
    """DeformConvV2 in training mode: loss.backward() fills every parameter gradient."""
    from detectron2_centernet_b200 import deform_conv as DC
    torch.manual_seed(0)
    layer = DC.DeformConvV2(64, 64).to(DEV).train()
    with torch.no_grad():
        layer.conv.conv_offset_mask.weight.normal_(0, 0.05)
    x = torch.randn(2, 64, 24, 24, device=DEV, requires_grad=True)
    layer(x).square().mean().backward()
    for n, prm in layer.named_parameters():
        assert prm.grad is not None and torch.isfinite(prm.grad).all(), n
    assert torch.isfinite(x.grad).all() and float(x.grad.abs().max()) > 0


def test_training_constraints_are_checked_in_forward():
    """Cout = 48 runs forward-only, but with autograd the unsupported backward shape is reported by the FORWARD call
    (it used to surface as a NotImplementedError inside autograd's backward)."""
    from detectron2_centernet_b200 import deform_conv as DC
    x = torch.randn(1, 64, 10, 11, device=DEV)
    off = torch.zeros(1, 18, 10, 11, device=DEV)
    m = torch.ones(1, 9, 10, 11, device=DEV)
    w = torch.randn(48, 64, 3, 3, device=DEV) * 0.05
    assert DC.modulated_deform_conv(x, off, m, w, None, 1, 1, 1, 1, 1).shape == (1, 48, 10, 11)
    with pytest.raises(NotImplementedError, match="out_channels"):
        DC.modulated_deform_conv(x.requires_grad_(True), off, m, w, None, 1, 1, 1, 1, 1)
