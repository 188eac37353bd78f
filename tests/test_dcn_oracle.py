"""CPU: the DCNv2 C oracle (oracle/dcn_oracle.c) against torchvision.ops.deform_conv2d(mask=...) and
against plain conv2d at zero offset / unit mask.  (The reference has no CPU implementation of this op
-- deform_conv.h:311 -- and its tests hold no vectors for it: parity is pinned on torchvision.)"""
import numpy as np
import pytest
import torch

from oracle import dcn_oracle as DO

tv = pytest.importorskip("torchvision.ops")


def _case(seed, B, Cin, Cout, H, W, s, p, d, off_range=3.0):
    rng = np.random.default_rng(seed)
    Ho = (H + 2 * p - (2 * d + 1)) // s + 1
    Wo = (W + 2 * p - (2 * d + 1)) // s + 1
    x = rng.normal(size=(B, Cin, H, W)).astype(np.float32)
    off = rng.uniform(-off_range, off_range, (B, 18, Ho, Wo)).astype(np.float32)
    m = rng.uniform(0, 1, (B, 9, Ho, Wo)).astype(np.float32)
    w = rng.normal(size=(Cout, Cin, 3, 3)).astype(np.float32)
    b = rng.normal(size=(Cout,)).astype(np.float32)
    return x, off, m, w, b


@pytest.mark.parametrize("cfg", [(2, 8, 16, 12, 10, 1, 1, 1), (1, 4, 8, 9, 11, 2, 1, 1), (1, 4, 8, 9, 11, 1, 2, 2),
                                 (1, 16, 4, 7, 7, 1, 0, 1)])
def test_oracle_matches_torchvision(cfg):
    B, Cin, Cout, H, W, s, p, d = cfg
    x, off, m, w, b = _case(1, B, Cin, Cout, H, W, s, p, d, off_range=6.0)    # many samples fall outside
    got = DO.dcn_forward(x, off, m, w, b, s, p, d, threads=2)
    ref = tv.deform_conv2d(torch.from_numpy(x), torch.from_numpy(off), torch.from_numpy(w), torch.from_numpy(b),
                           stride=s, padding=p, dilation=d, mask=torch.from_numpy(m)).numpy()
    np.testing.assert_allclose(got, ref, rtol=1e-4, atol=1e-4)


def test_zero_offset_unit_mask_is_conv2d():
    x, off, m, w, b = _case(2, 1, 6, 5, 10, 9, 1, 1, 1)
    got = DO.dcn_forward(x, np.zeros_like(off), np.ones_like(m), w, b, 1, 1, 1)
    ref = torch.nn.functional.conv2d(torch.from_numpy(x), torch.from_numpy(w), torch.from_numpy(b), padding=1).numpy()
    np.testing.assert_allclose(got, ref, rtol=1e-4, atol=1e-4)


def test_border_rule():
    """A tap at exactly h_im = -1 contributes nothing; just inside (-1, 0) blends with zero rows."""
    x = np.ones((1, 1, 4, 4), np.float32)
    w = np.zeros((1, 1, 3, 3), np.float32)
    w[0, 0, 1, 1] = 1.0                                   # centre tap only
    m = np.ones((1, 9, 4, 4), np.float32)
    off = np.zeros((1, 18, 4, 4), np.float32)
    off[0, 8] = -1.0                                      # tap 4 (centre): dh = -1 -> row 0 samples h_im = -1
    out = DO.dcn_forward(x, off, m, w, None, 1, 1, 1)
    assert np.all(out[0, 0, 0] == 0.0) and np.all(out[0, 0, 1:] == 1.0)
    off[0, 8] = -0.75                                     # row 0 samples h_im = -0.75: 0.25 * row 0
    out = DO.dcn_forward(x, off, m, w, None, 1, 1, 1)
    np.testing.assert_allclose(out[0, 0, 0], 0.25, rtol=1e-6)


@pytest.mark.parametrize("cfg", [(2, 8, 16, 12, 10, 1, 1, 1), (1, 4, 8, 9, 11, 2, 1, 1), (1, 4, 8, 9, 11, 1, 2, 2)])
def test_oracle_backward_matches_torchvision_autograd(cfg):
    """The backward restatement (col2im_coord / col2im / im2col + GEMMs) against autograd through
    torchvision's op; offsets reach far outside the image so the border rules are exercised."""
    B, Cin, Cout, H, W, s, p, d = cfg
    x, off, m, w, b = _case(5, B, Cin, Cout, H, W, s, p, d, off_range=5.0)
    rng = np.random.default_rng(9)
    Ho, Wo = off.shape[2], off.shape[3]
    gy = rng.normal(size=(B, Cout, Ho, Wo)).astype(np.float32)
    t = [torch.from_numpy(a).double().requires_grad_(True) for a in (x, off, m, w, b)]
    tv.deform_conv2d(t[0], t[1], t[3], t[4], stride=s, padding=p, dilation=d, mask=t[2]).backward(torch.from_numpy(gy).double())
    got = DO.dcn_backward(x, off, m, w, gy, s, p, d)
    for g, r, name in zip(got, (t[0], t[1], t[2], t[3], t[4]), ("input", "offset", "mask", "weight", "bias")):
        np.testing.assert_allclose(g, r.grad.numpy(), rtol=2e-3, atol=2e-3 * float(r.grad.abs().max()), err_msg=name)
