"""GPU: SURVEY section 8 f2, second half, decode side -- `ctdet_head_decode_fused` (the hm head's final 1x1 convolution +
sigmoid + ctdet_decode without the logits in HBM) against the reference chain
    Conv2d(head_conv, C, 1) (centernet.py:118-123, applied :152) -> hm.sigmoid_() (:161) -> ctdet_decode (:426-458).

Two-step parity, because the convolution's operands are rounded to bf16 for the tensor cores:
  1. the logits the kernel forms (its optional `logits_out`) against torch's fp32 conv2d: relative L2 <= 1e-2 (stated bf16
     tolerance; exact to fp32 summation order when feat / weight are bf16-representable: <= 1e-5);
  2. the detections against `ctdet_decode` (pinned on the reference's golden vectors in tests/test_gpu_parity.py) run on
     THOSE logits: bit for bit (`torch.equal`) -- boxes, scores, classes, inds -- with and without `logits_out`.
With bf16-representable, well-separated inputs the detections are also compared with the oracle port of the reference's own
ctdet_decode on torch's fp32 logits (classes / inds equal, scores and boxes to 1e-5).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(seed, B, Cf, C, H, W=128, bf16_exact=False, peaks=60):
    g = torch.Generator(device="cpu").manual_seed(seed)
    feat = torch.relu(torch.randn(B, Cf, H, W, generator=g))
    w = torch.randn(C, Cf, generator=g) * (1.0 / Cf ** 0.5)
    b = torch.full((C,), -2.19)
    # planted objects: a few cells whose features line up with one class's weights
    for bi in range(B):
        for _ in range(peaks):
            c = int(torch.randint(0, C, (1,), generator=g))
            y, x = int(torch.randint(0, H, (1,), generator=g)), int(torch.randint(0, W, (1,), generator=g))
            feat[bi, :, y, x] += torch.relu(w[c]) * float(torch.rand(1, generator=g) * 6 + 2) * Cf ** 0.5 / 8
    if bf16_exact:
        feat, w = feat.bfloat16().float(), w.bfloat16().float()
    wh = torch.rand(B, 2, H, W, generator=g) * 20 + 1
    reg = torch.rand(B, 2, H, W, generator=g)
    return [t.to(DEV) for t in (feat, w, b, wh, reg)]


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("cfg", [(2, 64, 16, 12, 40), (3, 256, 80, 32, 100), (2, 128, 32, 128, 100), (5, 256, 80, 64, 100),
                                 (1, 64, 16, 1, 20), (2, 64, 16, 2, 30), (1, 192, 48, 33, 100),
                                 (2, 256, 80, 104, 100, 104), (3, 64, 16, 40, 60, 40), (2, 64, 32, 9, 25, 7), (1, 128, 16, 50, 100, 97)])
def test_fused_head_decode_equals_decode_of_its_own_logits(cfg):
    """Shapes 5-7: one and two map rows (no halo rows exist), an odd row count (runs that end inside an image).  The last
    four are narrow maps (W < 128: a 416^2 input at stride 4 is 104 x 104; odd widths take ctdet_decode's own odd-width
    path in the check)."""
    from detectron2_centernet_b200 import decode as D
    B, Cf, C, H, K = cfg[:5]
    W = cfg[5] if len(cfg) > 5 else 128
    feat, w, b, wh, reg = _setup(3 + H, B, Cf, C, H, W)
    boxes, scores, classes, inds, logits = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K, return_logits=True)
    want_logits = torch.nn.functional.conv2d(feat, w[:, :, None, None], b)
    assert _rel(logits, want_logits) <= 1e-2
    want = D.ctdet_decode_batch(logits, wh, reg, down_ratio=4, K=K, logits=True)
    for name, a, r in zip(("boxes", "scores", "classes", "inds"), (boxes, scores, classes, inds), want):
        assert torch.equal(a, r), (cfg, name)
    again = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K)          # without the logits pass
    for name, a, r in zip(("boxes", "scores", "classes", "inds"), again, want):
        assert torch.equal(a, r), (cfg, name, "no logits_out")


def test_fused_head_decode_matches_reference_chain_on_bf16_exact_inputs():
    from oracle import ctdet_oracle as O
    from detectron2_centernet_b200 import decode as D
    B, Cf, C, H, K = 2, 128, 16, 24, 50
    feat, w, b, wh, reg = _setup(91, B, Cf, C, H, bf16_exact=True)
    boxes, scores, classes, inds, logits = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K, return_logits=True)
    want_logits = torch.nn.functional.conv2d(feat.double(), w.double()[:, :, None, None], b.double()).float()
    assert _rel(logits, want_logits) <= 1e-5
    rb, rs, rc, ri = O.decode(want_logits.cpu().numpy(), wh.cpu().numpy(), reg.cpu().numpy(), down_ratio=4, K=K, logits=True)
    for bi in range(B):
        # the top detections are planted, well-separated peaks: compare while the reference's scores are distinct by > 1e-4
        n = K
        for i in range(1, K):
            if rs[bi, i - 1] - rs[bi, i] < 1e-4:
                n = i
                break
        assert n >= 10
        assert np.array_equal(classes[bi, :n].cpu().numpy(), rc[bi, :n])
        assert np.array_equal(inds[bi, :n].cpu().numpy(), ri[bi, :n])
        np.testing.assert_allclose(scores[bi, :n].cpu().numpy(), rs[bi, :n], rtol=0, atol=2e-5)
        np.testing.assert_allclose(boxes[bi, :n].cpu().numpy(), rb[bi, :n], rtol=0, atol=2e-3)


def test_plateaus_take_the_exact_fallback():
    """weight = 0: every logit equals the bias, every cell ties with its neighbours (a peak by the reference's `hmax == heat`),
    every key segment overflows -> the images are handed to the staged path on the stored logits; one image of the batch is
    normal, so both paths run in one call."""
    from detectron2_centernet_b200 import decode as D
    B, Cf, C, H, K = 3, 64, 16, 16, 100
    feat, w, b, wh, reg = _setup(5, B, Cf, C, H)
    feat[0].zero_()
    feat[2].zero_()
    out = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K, return_logits=True)
    want = D.ctdet_decode_batch(out[4], wh, reg, down_ratio=4, K=K, logits=True)
    for name, a, r in zip(("boxes", "scores", "classes", "inds"), out[:4], want):
        assert torch.equal(a, r), name
    assert torch.equal(out[3][0], torch.arange(K, device=DEV, dtype=torch.int64))      # ties: lowest flat index first
    out2 = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K)               # fall-back on the workspace's logits
    for name, a, r in zip(("boxes", "scores", "classes", "inds"), out2, want):
        assert torch.equal(a, r), (name, "no logits_out")


def test_object_concentrated_scores_use_the_spill_list():
    """A hot region (many cells far above the rest in many classes): its rows' (row, class half) segments hold more than 64
    keys -> the surplus goes to the image's spill list; still bit-equal to ctdet_decode on the kernel's logits."""
    from detectron2_centernet_b200 import decode as D
    B, Cf, C, H, K = 2, 64, 32, 64, 100
    feat, w, b, wh, reg = _setup(31, B, Cf, C, H, peaks=0)
    g = torch.Generator(device="cpu").manual_seed(2)
    feat[1, :, 20:26, 10:90] *= 1.0 + 2.0 * torch.rand(Cf, 6, 80, generator=g).to(DEV)
    out = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K, return_logits=True)
    want = D.ctdet_decode_batch(out[4], wh, reg, down_ratio=4, K=K, logits=True)
    for name, a, r in zip(("boxes", "scores", "classes", "inds"), out[:4], want):
        assert torch.equal(a, r), name
    out2 = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K)
    for name, a, r in zip(("boxes", "scores", "classes", "inds"), out2, want):
        assert torch.equal(a, r), (name, "second call")


def test_two_calls_in_a_row_and_shape_changes_share_the_workspace():
    from detectron2_centernet_b200 import decode as D
    for i, (B, Cf, C, H, K) in enumerate([(2, 64, 16, 40, 100), (2, 64, 16, 40, 100), (2, 128, 32, 8, 50), (2, 64, 16, 40, 100)]):
        feat, w, b, wh, reg = _setup(20 + i, B, Cf, C, H)
        out = D.hm_head_decode_batch(feat, w, b, wh, reg, down_ratio=4, K=K, return_logits=True)
        want = D.ctdet_decode_batch(out[4], wh, reg, down_ratio=4, K=K, logits=True)
        for name, a, r in zip(("boxes", "scores", "classes", "inds"), out[:4], want):
            assert torch.equal(a, r), (i, name)


def test_unsupported_shapes_are_rejected():
    from detectron2_centernet_b200 import _lib
    from detectron2_centernet_b200 import decode as D
    lib = _lib.load()
    assert lib.ctdet_head_decode_fused_workspace_bytes(2, 64, 16, 16, 132, 10) == 0         # W > 128
    assert lib.ctdet_head_decode_fused_workspace_bytes(2, 60, 16, 16, 128, 10) == 0         # Cf % 64
    assert lib.ctdet_head_decode_fused_workspace_bytes(2, 64, 20, 16, 128, 10) == 0         # C % 16
    assert not D.head_decode_fusable(64, 16, 16, 256) and D.head_decode_fusable(64, 16, 16, 64)
    feat = torch.zeros(1, 64, 4, 256, device=DEV)
    with pytest.raises(NotImplementedError):
        D.hm_head_decode_batch(feat, torch.zeros(16, 64, device=DEV), None, torch.zeros(1, 2, 4, 256, device=DEV))


def test_ctdet_head_decode_method_fused_vs_unfused():
    """`CtdetHead.decode` (the eval path of centernet.py:151-166 on the head modules): the fused and the unfused route agree
    on the planted, well-separated top detections (classes / cells equal, scores to 2e-3: bf16 operands)."""
    from detectron2_centernet_b200 import neck as N
    torch.manual_seed(7)
    head = N.CtdetHead(64, 80, 256).to(DEV).eval()
    x = torch.randn(2, 64, 16, 128, device=DEV)
    with torch.no_grad():
        last = head.hm[-1]
        feat = head.hm[:-1](x)
        for b in range(2):                                     # plant 12 peaks per image through the last layer's weights
            for i in range(12):
                c, yy, xx = (7 * i + b) % 80, (5 * i + 1) % 16, (19 * i + 3) % 128
                # raise the 3x3-conv output so that class c's logit at the cell is high: solve roughly through the bias of hm[0]
                feat[b, :, yy, xx] += torch.relu(last.weight[c, :, 0, 0]) * (40.0 + 5 * i)
        wh, reg = head.wh(x), head.reg(x)
        from detectron2_centernet_b200 import decode as D
        fused = D.hm_head_decode_batch(feat, last.weight, last.bias, wh, reg, down_ratio=4, K=50)
        plain = D.ctdet_decode_batch(last(feat), wh, reg, down_ratio=4, K=50, logits=True)
        assert torch.equal(fused[2][:, :8], plain[2][:, :8]) and torch.equal(fused[3][:, :8], plain[3][:, :8])
        assert float((fused[1][:, :8] - plain[1][:, :8]).abs().max()) <= 2e-3
        a, b_ = head.decode(x, K=50), head.decode(x, K=50, fuse=False)
        assert a[1].shape == b_[1].shape == (2, 50)
        assert float((a[1][:, 0] - b_[1][:, 0]).abs().max()) <= 2e-3


def test_running_bound_variant_in_a_fresh_process():
    """`CTDET_HEAD_DECODE_SEED=0` (read once per process) selects the sweep with a running bound per CTA (no seed launch)
    instead of the default seed launch + sweep: the same parity tests in a child process."""
    import os
    import subprocess
    import sys
    if os.environ.get("CTDET_HEAD_DECODE_SEED"):
        pytest.skip("already the child")
    env = dict(os.environ, CTDET_HEAD_DECODE_SEED="0")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-k",
                        "own_logits or plateaus or spill or share_the_workspace"], env=env, capture_output=True, text=True,
                       timeout=300, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
