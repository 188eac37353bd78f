"""GPU: out-of-bounds canaries for the round-2 kernels (compute-sanitizer is not available on the GPU pool): every output
and workspace of `ctdet_head_focal_fused_bwd` and `dcnv2_backward` (GEMM-scatter dgrad, paired wgrad) is carved out of the
middle of a larger allocation whose guard bands are filled with a bit pattern; after the call through the C ABI the guards
must be untouched and the outputs fully written (no cell left at the pattern)."""
import ctypes
import math

import numpy as np
import pytest
import torch

from detectron2_centernet_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GUARD = 4096            # bytes either side
PATTERN = 0x7FC12345    # a quiet NaN with a payload


class Carved:
    """`nbytes` in the middle of a buffer with GUARD bytes of PATTERN either side (and inside, to detect unwritten cells)."""

    def __init__(self, shape, dtype):
        self.shape, self.dtype = tuple(shape), dtype
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        self.nbytes = (n + 255) // 256 * 256
        self.buf = torch.full(((2 * GUARD + self.nbytes) // 4,), PATTERN, dtype=torch.int32, device=DEV)
        self.used = n

    def ptr(self):
        return ctypes.c_void_p(self.buf.data_ptr() + GUARD)

    def view(self):
        flat = self.buf.view(torch.uint8)[GUARD:GUARD + self.used]
        return flat.view(self.dtype).view(self.shape)

    def guards_intact(self):
        head = self.buf[:GUARD // 4]
        tail = self.buf[(GUARD + self.nbytes) // 4:]
        pad = self.buf.view(torch.uint8)[GUARD + self.used:GUARD + self.nbytes]
        ok_pad = True
        if pad.numel() >= 4 and self.used % 4 == 0:
            ok_pad = bool((pad.view(torch.int32) == PATTERN).all())
        return bool((head == PATTERN).all()) and bool((tail == PATTERN).all()) and ok_pad

    def fully_written(self):
        v = self.view()
        if v.dtype == torch.float32:
            return not bool((v.view(torch.int32) == PATTERN).any())
        return True


def test_head_focal_fused_bwd_stays_inside_its_buffers():
    lib = _lib.load()
    B, Cf, C, H, W = 3, 256, 80, 16, 24                      # 9 tiles of 128 pixels
    g = torch.Generator(device="cpu").manual_seed(3)
    feat = torch.relu(torch.randn(B, Cf, H, W, generator=g)).to(DEV)
    w = (torch.randn(C, Cf, generator=g) / 16).to(DEV)
    b = torch.full((C,), -2.19, device=DEV)
    gt = torch.rand(B, C, H, W, generator=g).pow(8).to(DEV)
    gt[0, 3, 5, 7] = 1.0
    alpha = torch.ones(C, device=DEV)
    norm = torch.tensor([1.0, 1.0], device=DEV)
    loss, d_logits = Carved((1,), torch.float32), Carved((B, C, H, W), torch.float32)
    d_feat, d_w, d_b = Carved((B, Cf, H, W), torch.float32), Carved((C, Cf), torch.float32), Carved((C,), torch.float32)
    ws = torch.zeros(lib.ctdet_loss_workspace_bytes(), dtype=torch.uint8, device=DEV)
    with torch.cuda.device(DEV):
        rc = lib.ctdet_head_focal_fused_bwd(_lib.ptr(feat), _lib.ptr(w), _lib.ptr(b), _lib.ptr(gt), _lib.ptr(alpha), None, _lib.ptr(norm),
                                            B, Cf, C, H, W, 2, 1.0, loss.ptr(), d_logits.ptr(), d_feat.ptr(), d_w.ptr(), d_b.ptr(),
                                            _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(DEV))
    assert rc == 0
    torch.cuda.synchronize()
    for name, c in (("loss", loss), ("d_logits", d_logits), ("d_feat", d_feat), ("d_weight", d_w), ("d_bias", d_b)):
        assert c.guards_intact(), name
        assert c.fully_written(), name
    assert not bool(ws.any())                                # the workspace contract: left zeroed
    # forward-only entry point
    loss2, d2 = Carved((1,), torch.float32), Carved((B, C, H, W), torch.float32)
    with torch.cuda.device(DEV):
        rc = lib.ctdet_head_focal_fused(_lib.ptr(feat), _lib.ptr(w), _lib.ptr(b), _lib.ptr(gt), _lib.ptr(alpha), None, _lib.ptr(norm),
                                        B, Cf, C, H, W, 2, 1.0, loss2.ptr(), d2.ptr(), _lib.ptr(ws), ctypes.c_size_t(ws.numel()),
                                        _lib.stream_ptr(DEV))
    assert rc == 0
    torch.cuda.synchronize()
    assert loss2.guards_intact() and d2.guards_intact() and d2.fully_written()
    assert torch.equal(d2.view(), d_logits.view()) and torch.equal(loss2.view(), loss.view())


@pytest.mark.parametrize("cfg", [(2, 64, 64, 20, 28, 2.5), (1, 128, 128, 12, 40, 6.0), (1, 256, 256, 9, 11, 2.0)])
def test_dcn_backward_stays_inside_its_buffers(cfg):
    """GEMM-scatter dgrad + paired wgrad (first case), Cout = 128 (second, offsets beyond the patch), the round-1 dgrad
    (third: Cout = 256); ragged tiles everywhere."""
    lib = _lib.load()
    B, Cin, Cout, H, W, off_range = cfg
    g = torch.Generator(device="cpu").manual_seed(9)
    x = torch.randn(B, Cin, H, W, generator=g).to(DEV)
    off = ((torch.rand(B, 18, H, W, generator=g) * 2 - 1) * off_range).to(DEV)
    m = torch.sigmoid(torch.randn(B, 9, H, W, generator=g)).to(DEV)
    wt = ((torch.rand(Cout, Cin, 3, 3, generator=g) * 2 - 1) / math.sqrt(9 * Cin)).to(DEV)
    gy = torch.randn(B, Cout, H, W, generator=g).to(DEV)
    gi, go, gm = Carved((B, Cin, H, W), torch.float32), Carved((B, 18, H, W), torch.float32), Carved((B, 9, H, W), torch.float32)
    gw, gb = Carved((Cout, Cin, 3, 3), torch.float32), Carved((Cout,), torch.float32)
    n = lib.dcnv2_backward_workspace_bytes(B, Cin, H, W, Cout, H, W)
    ws = Carved((n,), torch.uint8)
    with torch.cuda.device(DEV):
        rc = lib.dcnv2_backward(_lib.ptr(x), _lib.ptr(wt), _lib.ptr(off), _lib.ptr(m), _lib.ptr(gy), gi.ptr(), gw.ptr(), gb.ptr(),
                                go.ptr(), gm.ptr(), B, Cin, H, W, Cout, 3, 3, 1, 1, 1, 1, 1, 1, 1, 1, 0,
                                ws.ptr(), ctypes.c_size_t(n), _lib.stream_ptr(DEV))
    assert rc == 0
    torch.cuda.synchronize()
    for name, c in (("grad_input", gi), ("grad_offset", go), ("grad_mask", gm), ("grad_weight", gw), ("grad_bias", gb), ("workspace", ws)):
        assert c.guards_intact(), (cfg, name)
        assert c.fully_written(), (cfg, name)
    from detectron2_centernet_b200 import deform_conv as DC
    want = DC.dcnv2_backward(x, off, m, wt, gy, 1, 1, 1)
    torch.cuda.synchronize()
    assert torch.equal(go.view(), want[1]) and torch.equal(gm.view(), want[2])          # deterministic outputs
    assert torch.allclose(gi.view(), want[0], rtol=1e-4, atol=1e-5)                      # fp32 atomics: order varies


def test_head_decode_fused_and_spilling_decode_stay_inside_their_buffers():
    """`ctdet_head_decode_fused` (workspace incl. the spill list and the fall-back logits, outputs, optional logits_out) and
    `ctdet_decode` on class-concentrated logits (tiles that spill): guards untouched, outputs fully written."""
    lib = _lib.load()
    B, Cf, C, H, W, K = 3, 64, 32, 40, 128, 100
    g = torch.Generator(device="cpu").manual_seed(9)
    feat = torch.relu(torch.randn(B, Cf, H, W, generator=g))
    feat[1, :, 10:16, 20:100] *= 3.0                          # a hot region: segments overflow into the spill list
    feat[2].zero_()                                           # a plateau: spill list overflows -> staged path on stored logits
    feat = feat.to(DEV)
    w = (torch.randn(C, Cf, generator=g) / 8).to(DEV)
    bias = torch.full((C,), -2.19, device=DEV)
    wh = (torch.rand(B, 2, H, W, generator=g) * 20).to(DEV)
    reg = torch.rand(B, 2, H, W, generator=g).to(DEV)
    for with_logits in (False, True):
        outs = [Carved((B, K, 4), torch.float32), Carved((B, K), torch.float32), Carved((B, K), torch.int32),
                Carved((B, K), torch.int64)]
        lg = Carved((B, C, H, W), torch.float32) if with_logits else None
        n = lib.ctdet_head_decode_fused_workspace_bytes(B, Cf, C, H, W, K)
        ws = Carved((n,), torch.uint8)
        ws.view().zero_()
        for _ in range(2):
            rc = lib.ctdet_head_decode_fused(_lib.ptr(feat), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(wh), _lib.ptr(reg),
                                             B, Cf, C, H, W, K, 4.0, 0, *[o.ptr() for o in outs], lg.ptr() if lg else None,
                                             ws.ptr(), ctypes.c_size_t(n), _lib.stream_ptr(DEV))
            assert rc == 0, lib.ctdet_error_string(rc)
            torch.cuda.synchronize()
        for o in outs + [ws] + ([lg] if lg else []):
            assert o.guards_intact()
        assert outs[0].fully_written() and outs[1].fully_written() and (lg is None or lg.fully_written())
        assert bool((outs[2].view() >= 0).all()) and bool((outs[2].view() < C).all())
        assert bool((outs[3].view() >= 0).all()) and bool((outs[3].view() < H * W).all())
    # the plain decode on logits with a per-class mean (tiles spill)
    C2, H2 = 80, 128
    hm = (0.7 * torch.randn(B, C2, H2, W, generator=g) + (-2.19 + 0.6 * torch.randn(1, C2, 1, 1, generator=g))).to(DEV)
    wh2 = (torch.rand(B, 2, H2, W, generator=g) * 20).to(DEV)
    outs = [Carved((B, K, 4), torch.float32), Carved((B, K), torch.float32), Carved((B, K), torch.int32),
            Carved((B, K), torch.int64)]
    n = lib.ctdet_decode_workspace_bytes(B, C2, H2, W, K)
    ws = Carved((n,), torch.uint8)
    ws.view().zero_()
    for _ in range(2):
        rc = lib.ctdet_decode(_lib.ptr(hm), _lib.ptr(wh2), None, B, C2, H2, W, K, 4.0, 1, *[o.ptr() for o in outs],
                              ws.ptr(), ctypes.c_size_t(n), _lib.stream_ptr(DEV))
        assert rc == 0, lib.ctdet_error_string(rc)
        torch.cuda.synchronize()
    for o in outs + [ws]:
        assert o.guards_intact()
    assert outs[0].fully_written() and outs[1].fully_written()
