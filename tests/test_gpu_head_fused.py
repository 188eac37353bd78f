"""GPU: SURVEY section 8 f2, second half -- `ctdet_head_focal_fused` (the hm head's final 1x1 convolution + focal loss in
one kernel; the logits never reach HBM) against the reference chain
    Conv2d(head_conv, C, 1) (centernet.py:118-123, applied :152) -> sigmoid_ + clamp (:204) -> _neg_loss (:333-369)
evaluated in fp32 by torch + the pinned oracle port (oracle/ref_gpu_chain.neg_loss), with autograd for the gradients.

Stated tolerance (feat and weight are rounded to bf16 for the tensor cores, fp32 accumulation over Cf = 256):
  loss <= 2e-3 relative;  d_logits relative L2 <= 2e-2, max <= 5e-2 of the max;  the convolution's parameter / input
  gradients (torch GEMMs on d_logits) relative L2 <= 2e-2.
With bf16-representable feat / weight the only difference to `ctdet_loss_fwd_bwd` on torch's fp32 logits is the summation
order of the 256-term dot products: d_logits relative L2 <= 1e-4.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(seed, B, Cf, C, H, W, bf16_exact=False):
    from detectron2_centernet_b200 import synthetic as S
    from detectron2_centernet_b200 import target as T
    g = torch.Generator(device="cpu").manual_seed(seed)
    feat = torch.relu(torch.randn(B, Cf, H, W, generator=g))              # the head's ReLU output
    w = torch.randn(C, Cf, generator=g) * (1.0 / Cf ** 0.5)
    b = torch.full((C,), -2.19)
    if bf16_exact:
        feat, w = feat.bfloat16().float(), w.bfloat16().float()
    bl, cl = S.make_batch_objects(seed, B, 4 * max(H, W), num_classes=C)
    boxes, classes, offsets = S.pack_objects(bl, cl)
    tg = T.target_gen(*[torch.from_numpy(np.ascontiguousarray(a)).to(DEV) for a in (boxes, classes, offsets)], B, C, H, W)
    return feat.to(DEV), w.to(DEV), b.to(DEV), tg


def _reference(feat, w, b, gt, alpha):
    from oracle import ref_gpu_chain as RC
    feat = feat.clone().requires_grad_(True)
    w = w.clone().requires_grad_(True)
    b = b.clone().requires_grad_(True)
    logits = torch.nn.functional.conv2d(feat, w[:, :, None, None], b)
    logits.retain_grad()
    hm = torch.clamp(torch.sigmoid(logits), min=1e-4, max=1 - 1e-4)
    loss = RC.neg_loss(hm, gt, alpha)
    loss.backward()
    return loss.detach(), logits.grad, feat.grad, w.grad, b.grad


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30)), float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("fuse_backward", [False, True])
@pytest.mark.parametrize("cfg", [(2, 256, 80, 32, 32), (1, 64, 16, 16, 8), (3, 128, 96, 24, 16), (5, 256, 80, 64, 64), (3, 256, 80, 13, 11),
                                 (2, 128, 32, 26, 26)])
def test_fused_head_loss_matches_conv_plus_reference_loss(cfg, fuse_backward):
    """fuse_backward: the convolution's backward inside the kernel where the shapes allow it (cases 1, 3 and 4: the last
    one runs several tiles per CTA), else d_logits + torch GEMMs.  The last two shapes have H*W % 128 != 0: the last tile of
    every image is ragged (143 = 128 + 15 pixels; 676 = 5 x 128 + 36 -- what a 416^2 input gives at stride 4 is 104^2 = 84.5 tiles)."""
    from detectron2_centernet_b200 import losses as L
    B, Cf, C, H, W = cfg
    feat, w, b, tg = _setup(11 + Cf, B, Cf, C, H, W)
    alpha = [1.0] if C != 16 else [0.5 + 0.1 * i for i in range(16)]      # per-class alpha on the small case
    want = _reference(feat, w, b, tg["hm"], alpha)
    f = feat.clone().requires_grad_(True)
    ww = w.clone().requires_grad_(True)
    bb = b.clone().requires_grad_(True)
    loss = L.hm_head_focal_loss(f, ww, bb, tg["hm"], alpha=alpha, num_pos=tg.get("num_pos"), fuse_backward=fuse_backward)
    loss.sum().backward()
    torch.cuda.synchronize()
    assert abs(float(loss) - float(want[0])) <= 2e-3 * abs(float(want[0])), (float(loss), float(want[0]))
    for name, got, ref in (("feat", f.grad, want[2]), ("weight", ww.grad, want[3]), ("bias", bb.grad, want[4])):
        l2, mx = _rel(got, ref)
        assert l2 <= 2e-2, (cfg, name, l2, mx)
    # the kernel's own output, d(loss)/d(logits)
    loss2, d = L._head_focal_call(feat, w, b, tg["hm"], L.alpha_tensor(alpha, C, feat.device), tg.get("num_pos"), None,
                                  L.alpha_is_uniform(alpha), 1.0, True)
    l2, mx = _rel(d, want[1])
    assert l2 <= 2e-2 and mx <= 5e-2, (cfg, "d_logits", l2, mx)
    assert float(loss2) == float(loss)
    if fuse_backward and L.head_bwd_fusable(Cf, C, H * W):           # the kernel's optional d_logits output
        out = L._head_focal_bwd_call(feat, w, b, tg["hm"], L.alpha_tensor(alpha, C, feat.device), tg.get("num_pos"), None,
                                     L.alpha_is_uniform(alpha), 1.0, want_d_logits=True)
        assert torch.equal(out[4], d) and float(out[0]) == float(loss)
        assert torch.equal(out[1], f.grad)
        assert torch.allclose(out[3], bb.grad, rtol=1e-5, atol=1e-7)      # fp32 atomics: the order varies from run to run


def test_fused_head_loss_equals_the_unfused_kernel_on_bf16_representable_operands():
    from detectron2_centernet_b200 import losses as L
    B, Cf, C, H, W = 2, 256, 80, 32, 32
    feat, w, b, tg = _setup(5, B, Cf, C, H, W, bf16_exact=True)
    logits = torch.nn.functional.conv2d(feat.double(), w.double()[:, :, None, None], b.double()).float()
    a = L.alpha_tensor(None, C, feat.device)
    losses, d_hm, _, _ = L._call(logits, None, None, tg["hm"], None, None, None, None, a, tg.get("num_pos"), None,
                                 L._FLAG_ALPHA_UNIFORM, (1.0, 0.1, 1.0), True)
    loss, d = L._head_focal_call(feat, w, b, tg["hm"], a, tg.get("num_pos"), None, True, 1.0, True)
    torch.cuda.synchronize()
    l2, mx = _rel(d, d_hm)
    assert l2 <= 1e-4, (l2, mx)
    assert abs(float(loss) - float(losses[0])) <= 1e-5 * abs(float(losses[0]))
    # counted normaliser (no num_pos given) and the forward-only call
    loss3, none = L._head_focal_call(feat, w, b, tg["hm"], a, None, None, True, 1.0, False)
    assert none is None and abs(float(loss3) - float(loss)) <= 1e-6 * abs(float(loss))


def test_fused_head_loss_rejects_what_it_does_not_tile():
    from detectron2_centernet_b200 import losses as L
    with pytest.raises(NotImplementedError):                                 # C % 16 != 0
        L.hm_head_focal_loss(torch.zeros(1, 64, 16, 8, device=DEV), torch.zeros(20, 64, device=DEV), None,
                             torch.zeros(1, 20, 16, 8, device=DEV))
    with pytest.raises(NotImplementedError):                                 # 3x3 kernel
        L.hm_head_focal_loss(torch.zeros(1, 64, 16, 8, device=DEV), torch.zeros(16, 64, 3, 3, device=DEV), None,
                             torch.zeros(1, 16, 16, 8, device=DEV))


def test_losses_dict_with_the_fused_head_equals_ctdet_losses_on_materialised_logits():
    """`ctdet_losses_fused_head` (CenterNet.losses, centernet.py:191-212, with the hm branch's last convolution folded in):
    wh_loss / off_loss and their gradients are the same kernel's -> bit-exact; hm_loss within the bf16 tolerance above."""
    from detectron2_centernet_b200 import losses as L
    B, Cf, C, H, W = 3, 256, 80, 32, 32
    feat, w, b, tg = _setup(77, B, Cf, C, H, W)
    g = torch.Generator(device="cpu").manual_seed(5)
    wh0 = (torch.randn(B, 2, H, W, generator=g) * 4).to(DEV)
    reg0 = torch.rand(B, 2, H, W, generator=g).to(DEV)

    def leaves():
        return [t.clone().requires_grad_(True) for t in (feat, w, b, wh0, reg0)]

    f1, w1, b1, wh1, reg1 = leaves()
    logits = torch.nn.functional.conv2d(f1, w1[:, :, None, None], b1)
    want = L.ctdet_losses({"hm": logits, "wh": wh1, "reg": reg1}, tg)
    sum(want.values()).backward()
    f2, w2, b2, wh2, reg2 = leaves()
    got = L.ctdet_losses_fused_head(f2, w2, b2, {"wh": wh2, "reg": reg2}, tg)
    sum(got.values()).backward()
    torch.cuda.synchronize()
    assert set(got) == {"hm_loss", "wh_loss", "off_loss"}
    assert float(got["wh_loss"]) == float(want["wh_loss"]) and float(got["off_loss"]) == float(want["off_loss"])
    assert torch.equal(wh2.grad, wh1.grad) and torch.equal(reg2.grad, reg1.grad)
    assert abs(float(got["hm_loss"]) - float(want["hm_loss"])) <= 2e-3 * abs(float(want["hm_loss"]))
    for name, a, r in (("feat", f2.grad, f1.grad), ("weight", w2.grad, w1.grad), ("bias", b2.grad, b1.grad)):
        assert _rel(a, r)[0] <= 2e-2, name
    with pytest.raises(RuntimeError, match="forms the hm logits itself"):
        L.ctdet_losses_fused_head(f2, w2, b2, {"hm": logits.detach(), "wh": wh2, "reg": reg2}, tg)
