"""GPU: the CUDA path (through the C ABI) against the oracle and the committed golden vectors.

Bars (BASELINE.json north_star): decode indices/classes and ind/reg_mask targets bit-exact on
tie-free inputs; heatmap targets within 1e-6 absolute; loss values and gradients within 1e-5
relative; boxes bit-exact in fp32 (stated tolerance: 0).
"""
import numpy as np
import pytest
import torch

from oracle import ctdet_oracle as O
from oracle import gen_golden as G
from detectron2_centernet_b200 import synthetic as S

pytestmark = pytest.mark.gpu

CASES = list(G.CASES)
DEV = "cuda:0"


def _cuda(*arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).to(DEV) for a in arrs]


def _targets(bl, cl, H, W, C):
    from detectron2_centernet_b200 import target as T
    boxes, classes, offsets = S.pack_objects(bl, cl)
    if len(classes) == 0:
        boxes, classes = np.zeros((1, 4), np.float32), np.zeros((1,), np.int64)
    b, c, o = _cuda(boxes, classes, offsets)
    return T.target_gen(b, c, o, len(bl), C, H, W)


@pytest.fixture(scope="module")
def inputs():
    cache = {}

    def get(case):
        if case not in cache:
            cache[case] = G.build_inputs(case)
        return cache[case]
    return get


# ------------------------------------------------------------------ target generation
@pytest.mark.parametrize("case", CASES)
def test_target_gen(case, golden, inputs):
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs(case)
    out = _targets(bl, cl, H, W, C)
    torch.cuda.synchronize()
    ref = O.gen_heatmap_batch(bl, cl, H, W, C)
    for k in ("reg_mask", "ind", "wh", "reg"):
        got = out[k].cpu().numpy()
        assert got.dtype == ref[k].dtype
        assert np.array_equal(got, ref[k]), k                       # bit-exact slots
        assert np.array_equal(got, golden[case + "/gt_" + k]), k    # and equal to the reference's
    got_hm = out["hm"].cpu().numpy()
    assert np.abs(got_hm - ref["hm"]).max() <= 1e-6                 # north_star: 1e-6 absolute
    assert np.array_equal(got_hm == 1, ref["hm"] == 1)              # positives identical
    assert np.array_equal(got_hm == 0, ref["hm"] == 0)              # footprints identical
    npos = out["num_pos"].cpu().numpy()
    assert npos.sum() == int(golden[case + "/num_pos"])
    assert np.array_equal(npos, (ref["hm"] == 1).reshape(B, -1).sum(1))


def test_target_gen_edge_cases():
    C, H, W = 3, 40, 52          # W % 4 == 0 (bulk store) but not a power of two
    bl = [np.array([[0, 0, 4, 4], [200, 150, 208, 160], [10, 10, 10, 50], [-30, -30, 10, 10],
                    [100, 60, 140, 100], [100, 60, 140, 100], [0, 0, 208, 160]], np.float32),
          np.zeros((0, 4), np.float32),
          np.tile(np.array([[40, 40, 80, 90]], np.float32), (200, 1))]      # 200 identical objects
    cl = [np.array([0, 1, 2, 0, 1, 1, 2]), np.zeros((0,), np.int64), np.arange(200) % C]
    out = _targets(bl, cl, H, W, C)
    ref = O.gen_heatmap_batch(bl, cl, H, W, C)
    for k in ("reg_mask", "ind", "wh", "reg"):
        assert np.array_equal(out[k].cpu().numpy(), ref[k]), k
    assert np.abs(out["hm"].cpu().numpy() - ref["hm"]).max() <= 1e-6
    assert np.array_equal(out["num_pos"].cpu().numpy(), (ref["hm"] == 1).reshape(3, -1).sum(1))


def test_target_gen_odd_width_fallback():
    C, H, W = 2, 9, 7            # W % 4 != 0: plain-store path
    bl, cl = S.make_batch_objects(5, 2, 28, n_max=5, num_classes=C)
    out = _targets(bl, cl, H, W, C)
    ref = O.gen_heatmap_batch(bl, cl, H, W, C)
    assert np.abs(out["hm"].cpu().numpy() - ref["hm"]).max() <= 1e-6
    assert np.array_equal(out["ind"].cpu().numpy(), ref["ind"])


def test_gen_heatmap_dropin_signature():
    """gen_heatmap(instances, output_shape, meta) as called at centernet.py:188."""
    from detectron2_centernet_b200.structures import Boxes, Instances
    from detectron2_centernet_b200.target import gen_heatmap
    bl, cl = S.make_batch_objects(9, 1, 512)
    inst = Instances((512, 512))
    inst.gt_boxes = Boxes(torch.from_numpy(bl[0]).to(DEV))
    inst.gt_classes = torch.from_numpy(cl[0]).to(DEV)

    class Meta:
        thing_classes = ["c%d" % i for i in range(80)]
    out = gen_heatmap(inst, np.array([128, 128]), Meta())
    ref = O.gen_heatmap(bl[0], cl[0], 128, 128, 80)
    assert out["hm"].shape == (80, 128, 128) and out["reg_mask"].dtype == torch.uint8
    assert out["ind"].dtype == torch.int64 and out["wh"].shape == (128, 2)
    assert np.abs(out["hm"].cpu().numpy() - ref["hm"]).max() <= 1e-6
    assert np.array_equal(out["ind"].cpu().numpy(), ref["ind"])


# ------------------------------------------------------------------ losses
def _loss_case(case, inputs, alpha, golden=None, tag=None):
    from detectron2_centernet_b200 import losses as L
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs(case)
    gt_np = O.gen_heatmap_batch(bl, cl, H, W, C)
    gt = {k: torch.from_numpy(v).to(DEV) for k, v in gt_np.items()}
    t_hm, t_wh, t_reg = [t.requires_grad_(True) for t in _cuda(hm, wh, reg)]
    return (bl, cl, hm, wh, reg, gt_np, gt, t_hm, t_wh, t_reg, L)


def _check_grads(got, ref, what):
    ref = ref.numpy() if isinstance(ref, torch.Tensor) else ref
    got = got.cpu().numpy()
    scale = np.abs(ref).max()
    # 1e-5 relative; the absolute floor (1e-6 of the largest gradient) covers cells whose fp32
    # reference value is itself dominated by rounding of log(1-p) for p ~ 1e-4
    np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-6 * scale, err_msg=what)


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("tag", ["a1", "acls"])
def test_fused_losses(case, tag, golden, inputs):
    bl, cl, hm, wh, reg, gt_np, gt, t_hm, t_wh, t_reg, L = _loss_case(case, inputs, None)
    C = hm.shape[1]
    alpha = [1.0] if tag == "a1" else [0.25 + 0.01 * c for c in range(C)]
    out = L.ctdet_losses({"hm": t_hm, "wh": t_wh, "reg": t_reg}, gt, alpha=alpha)
    (out["hm_loss"] + out["wh_loss"] + out["off_loss"]).backward()
    torch.cuda.synchronize()
    got = np.array([out[k].item() for k in ("hm_loss", "wh_loss", "off_loss")])
    np.testing.assert_allclose(got, golden["%s/loss_%s" % (case, tag)], rtol=1e-5)   # vs the reference
    lo, gr = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg),
                                gt_np, alpha=alpha)
    np.testing.assert_allclose(got, [float(lo[k]) for k in ("hm_loss", "wh_loss", "off_loss")], rtol=1e-5)
    _check_grads(t_hm.grad, gr[0], "d_hm")
    _check_grads(t_wh.grad, gr[1], "d_wh")
    _check_grads(t_reg.grad, gr[2], "d_reg")
    # hm logits are NOT modified in place (the reference's sigmoid_ at :204 is)
    assert np.array_equal(t_hm.detach().cpu().numpy(), hm)


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("tag", ["a1", "acls"])
def test_fused_targets_and_losses_from_objects(case, tag, golden, inputs):
    """ctdet_target_loss_fused (gen_heatmap + CenterNet.losses in one pass, the heatmap target never
    written to HBM): same losses / gradients / slot targets as the reference chain."""
    from detectron2_centernet_b200 import losses as L
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs(case)
    alpha = [1.0] if tag == "a1" else [0.25 + 0.01 * c for c in range(C)]
    boxes, classes, offsets = S.pack_objects(bl, cl)
    if len(classes) == 0:
        boxes, classes = np.zeros((1, 4), np.float32), np.zeros((1,), np.int64)
    t_b, t_c, t_o = _cuda(boxes, classes, offsets)
    t_hm, t_wh, t_reg = [t.requires_grad_(True) for t in _cuda(hm, wh, reg)]
    out, slots = L.ctdet_losses_from_objects({"hm": t_hm, "wh": t_wh, "reg": t_reg}, t_b, t_c, t_o, alpha=alpha)
    (out["hm_loss"] + out["wh_loss"] + out["off_loss"]).backward()
    torch.cuda.synchronize()
    got = np.array([out[k].item() for k in ("hm_loss", "wh_loss", "off_loss")])
    np.testing.assert_allclose(got, golden["%s/loss_%s" % (case, tag)], rtol=1e-5)   # vs the reference
    gt_np = O.gen_heatmap_batch(bl, cl, H, W, C)
    lo, gr = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), gt_np, alpha=alpha)
    _check_grads(t_hm.grad, gr[0], "d_hm")
    _check_grads(t_wh.grad, gr[1], "d_wh")
    _check_grads(t_reg.grad, gr[2], "d_reg")
    for k in ("reg_mask", "ind", "wh", "reg"):
        assert np.array_equal(slots[k].cpu().numpy(), gt_np[k]), k
    assert int(slots["num_pos"].sum().item()) == int((gt_np["hm"] == 1).sum())


@pytest.mark.parametrize("shape", [(2, 3, 40, 36), (1, 5, 129, 132), (3, 2, 8, 4), (2, 4, 33, 256)])
def test_fused_from_objects_edge_shapes(shape):
    """ctdet_target_loss_fused on ragged / awkward inputs: maps that are not a multiple of the band
    height, a partial last band, several bands per plane, images without objects, objects stacked
    on one centre, boxes hanging over the border, saturated logits (clamp active -> zero gradient).
    Checked against the oracle chain gen_heatmap -> CenterNet.losses on the same inputs."""
    from detectron2_centernet_b200 import losses as L
    B, C, H, W = shape
    rng = np.random.default_rng(H * 1000 + W)
    size_y, size_x = 4 * H, 4 * W
    bl, cl = [], []
    for b in range(B):
        n = 0 if b == B - 1 and B > 1 else int(rng.integers(1, 12))
        # centres inside the map (the reference's gather raises otherwise), boxes may hang over the border
        cx = rng.uniform(0.5, size_x - 0.5, n); cy = rng.uniform(0.5, size_y - 0.5, n)
        bw = rng.uniform(2, size_x / 2, n); bh = rng.uniform(2, size_y / 2, n)
        boxes = np.stack([cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2], 1).astype(np.float32)
        classes = rng.integers(0, C, n).astype(np.int64)
        if n >= 3:                                   # two objects of one class on the same centre, different sizes
            boxes[1] = boxes[0] + np.array([-6, -4, 6, 4], np.float32)
            classes[1] = classes[0]
        bl.append(boxes); cl.append(classes)
    hm = rng.normal(-2.19, 1.5, (B, C, H, W)).astype(np.float32)
    hm.reshape(-1)[:: max(1, hm.size // 37)] = 14.0     # clamp active at 1-1e-4
    hm.reshape(-1)[1:: max(1, hm.size // 31)] = -15.0   # clamp active at 1e-4
    wh = rng.uniform(0, 64, (B, 2, H, W)).astype(np.float32)
    reg = rng.uniform(0, 1, (B, 2, H, W)).astype(np.float32)
    alpha = [0.3 + 0.05 * c for c in range(C)]
    boxes, classes, offsets = S.pack_objects(bl, cl)
    t_b, t_c, t_o = _cuda(boxes, classes, offsets)
    t_hm, t_wh, t_reg = [t.requires_grad_(True) for t in _cuda(hm, wh, reg)]
    out, slots = L.ctdet_losses_from_objects({"hm": t_hm, "wh": t_wh, "reg": t_reg}, t_b, t_c, t_o, alpha=alpha)
    (out["hm_loss"] + out["wh_loss"] + out["off_loss"]).backward()
    torch.cuda.synchronize()
    gt_np = O.gen_heatmap_batch(bl, cl, H, W, C)
    lo, gr = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), gt_np, alpha=alpha)
    got = np.array([out[k].item() for k in ("hm_loss", "wh_loss", "off_loss")])
    np.testing.assert_allclose(got, [float(lo[k]) for k in ("hm_loss", "wh_loss", "off_loss")], rtol=1e-5)
    _check_grads(t_hm.grad, gr[0], "d_hm")
    _check_grads(t_wh.grad, gr[1], "d_wh")
    _check_grads(t_reg.grad, gr[2], "d_reg")
    sat = (hm == 14.0) | (hm == -15.0)
    assert float(np.abs(t_hm.grad.cpu().numpy()[sat & (gt_np["hm"] < 1)]).max()) == 0.0     # torch.clamp passes no gradient
    for k in ("reg_mask", "ind", "wh", "reg"):
        assert np.array_equal(slots[k].cpu().numpy(), gt_np[k]), k
    # and the two-pass form gives the same numbers (same device functions)
    gt = _targets(bl, cl, H, W, C)
    u_hm = torch.from_numpy(hm).to(DEV).requires_grad_(True)
    out2 = L.ctdet_losses({"hm": u_hm, "wh": torch.from_numpy(wh).to(DEV), "reg": torch.from_numpy(reg).to(DEV)}, gt, alpha=alpha)
    out2["hm_loss"].backward()
    np.testing.assert_allclose(out["hm_loss"].item(), out2["hm_loss"].item(), rtol=1e-6)
    np.testing.assert_allclose(t_hm.grad.cpu().numpy(), u_hm.grad.cpu().numpy(), rtol=1e-6, atol=0)


def test_fused_losses_with_num_pos_from_target_gen(inputs):
    """num_pos produced by ctdet_target_gen feeds the loss: no counting pre-pass, same numbers."""
    from detectron2_centernet_b200 import losses as L
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs("coco512")
    gt = _targets(bl, cl, H, W, C)
    assert "num_pos" in gt
    t_hm, t_wh, t_reg = [t.requires_grad_(True) for t in _cuda(hm, wh, reg)]
    out = L.ctdet_losses({"hm": t_hm, "wh": t_wh, "reg": t_reg}, gt, alpha=[1.0])
    sum(out.values()).backward()
    gt2 = {k: v for k, v in gt.items() if k != "num_pos"}
    u_hm, u_wh, u_reg = [t.requires_grad_(True) for t in _cuda(hm, wh, reg)]
    out2 = L.ctdet_losses({"hm": u_hm, "wh": u_wh, "reg": u_reg}, gt2, alpha=[1.0])
    sum(out2.values()).backward()
    for k in out:
        np.testing.assert_allclose(out[k].item(), out2[k].item(), rtol=1e-6)
    np.testing.assert_allclose(t_hm.grad.cpu().numpy(), u_hm.grad.cpu().numpy(), rtol=1e-6, atol=0)


def test_losses_no_positives_and_upstream_grad(inputs):
    """num_pos == 0 branch (:365-366) and a non-unit upstream gradient."""
    from detectron2_centernet_b200 import losses as L
    B, C, H, W = 2, 3, 16, 16
    hm, wh, reg = S.make_head_outputs(3, B, C, H, W)
    gt_np = O.gen_heatmap_batch([np.zeros((0, 4), np.float32)] * B, [np.zeros((0,), np.int64)] * B, H, W, C)
    gt = {k: torch.from_numpy(v).to(DEV) for k, v in gt_np.items()}
    t_hm, t_wh, t_reg = [t.requires_grad_(True) for t in _cuda(hm, wh, reg)]
    out = L.ctdet_losses({"hm": t_hm, "wh": t_wh, "reg": t_reg}, gt, alpha=[1.0])
    (out["hm_loss"] * 3.0 + out["wh_loss"] * 0.5 + out["off_loss"]).backward()
    lo, _ = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), gt_np)
    x = torch.from_numpy(hm).requires_grad_(True)
    w = torch.from_numpy(wh).requires_grad_(True)
    r = torch.from_numpy(reg).requires_grad_(True)
    l2 = O.losses(x, w, r, gt_np)
    (l2["hm_loss"] * 3.0 + l2["wh_loss"] * 0.5 + l2["off_loss"]).backward()
    np.testing.assert_allclose(out["hm_loss"].item(), float(lo["hm_loss"]), rtol=1e-5)
    assert out["wh_loss"].item() == 0.0 and out["off_loss"].item() == 0.0
    _check_grads(t_hm.grad, x.grad, "d_hm (x3)")
    assert t_wh.grad.abs().max().item() == 0.0


def test_focal_and_regl1_dropins(inputs):
    """FocalLoss(alpha)(pred, gt) on probabilities and RegL1Loss()(output, mask, ind, target)."""
    from detectron2_centernet_b200 import losses as L
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs("dense_small")
    gt_np = O.gen_heatmap_batch(bl, cl, H, W, C)
    pred_np = S.sigmoid_clamp(hm)
    pred = torch.from_numpy(pred_np).to(DEV).requires_grad_(True)
    loss = L.FocalLoss([0.5])(pred, torch.from_numpy(gt_np["hm"]).to(DEV))
    loss.backward()
    p_ref = torch.from_numpy(pred_np).requires_grad_(True)
    l_ref = O.focal_loss(p_ref, torch.from_numpy(gt_np["hm"]), [0.5])
    l_ref.backward()
    np.testing.assert_allclose(loss.item(), l_ref.item(), rtol=1e-5)
    _check_grads(pred.grad, p_ref.grad, "d_pred")
    o = torch.from_numpy(wh).to(DEV).requires_grad_(True)
    l1 = L.RegL1Loss()(o, torch.from_numpy(gt_np["reg_mask"]).to(DEV),
                       torch.from_numpy(gt_np["ind"]).to(DEV), torch.from_numpy(gt_np["wh"]).to(DEV))
    l1.backward()
    o_ref = torch.from_numpy(wh).requires_grad_(True)
    l1_ref = O.reg_l1_loss(o_ref, torch.from_numpy(gt_np["reg_mask"]), torch.from_numpy(gt_np["ind"]),
                           torch.from_numpy(gt_np["wh"]))
    l1_ref.backward()
    np.testing.assert_allclose(l1.item(), l1_ref.item(), rtol=1e-5)
    _check_grads(o.grad, o_ref.grad, "d_output")


# ------------------------------------------------------------------ decode
@pytest.mark.parametrize("case", CASES)
def test_decode(case, golden, inputs):
    from detectron2_centernet_b200 import decode as D
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs(case)
    heat = S.sigmoid_clamp(hm)
    t_heat, t_wh, t_reg = _cuda(heat, wh, reg)
    boxes, scores, classes, inds = D.ctdet_decode_batch(t_heat, t_wh, t_reg, down_ratio=4, K=K)
    torch.cuda.synchronize()
    ob, os_, oc, oi = O.decode(heat, wh, reg, 4, K)
    assert classes.dtype == torch.int32 and inds.dtype == torch.int64
    assert np.array_equal(classes.cpu().numpy(), oc)
    assert np.array_equal(inds.cpu().numpy(), oi)
    assert np.array_equal(scores.cpu().numpy(), os_)
    assert np.array_equal(boxes.cpu().numpy(), ob)                       # fp32 bit-exact
    # and identical to what the reference itself produced
    assert np.array_equal(classes.cpu().numpy(), golden[case + "/dec_classes"])
    assert np.array_equal(scores.cpu().numpy(), golden[case + "/dec_scores"])
    assert np.array_equal(boxes.cpu().numpy(), golden[case + "/dec_boxes"])
    b2, _, _, _ = D.ctdet_decode_batch(t_heat, t_wh, None, down_ratio=4, K=K)
    assert np.array_equal(b2.cpu().numpy(), golden[case + "/dec_boxes_noreg"])
    # logits mode: same ranking, scores within 1 ulp of clamp(sigmoid)
    (t_hm,) = _cuda(hm)
    b3, s3, c3, i3 = D.ctdet_decode_batch(t_hm, t_wh, t_reg, down_ratio=4, K=K, logits=True)
    assert np.array_equal(c3.cpu().numpy(), oc) and np.array_equal(i3.cpu().numpy(), oi)
    assert np.array_equal(b3.cpu().numpy(), ob)
    np.testing.assert_allclose(s3.cpu().numpy(), os_, rtol=2e-7)
    # single-image drop-in signature
    bb, ss, cc = D.ctdet_decode(t_heat[:1], t_wh[:1], reg=t_reg[:1], down_ratio=4, K=K)
    assert bb.shape == (K, 4) and ss.shape == (K,) and cc.shape == (K,)
    assert np.array_equal(cc.cpu().numpy(), oc[0])


def test_decode_ties_plateaus_and_few_peaks():
    from detectron2_centernet_b200 import decode as D
    rng = np.random.default_rng(0)
    # (1) constant map: every cell is a peak with the same score -> lowest flat indices win
    heat = np.full((2, 3, 16, 20), 0.25, np.float32)
    # (2) second image: quantised values -> many exact ties and plateaus
    heat[1] = (rng.integers(1, 6, (3, 16, 20)) / 8.0).astype(np.float32)
    wh = rng.uniform(0, 8, (2, 2, 16, 20)).astype(np.float32)
    reg = rng.uniform(0, 1, (2, 2, 16, 20)).astype(np.float32)
    out = D.ctdet_decode_batch(*_cuda(heat, wh, reg), down_ratio=4, K=50)
    ref = O.decode(heat, wh, reg, 4, 50)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)
    # (3) fewer peaks than K: zeros (heat*keep) fill the tail in flat-index order
    # a strictly increasing ramp has exactly one peak per plane (its last cell)
    heat = (np.arange(2 * 64, dtype=np.float32).reshape(1, 2, 8, 8) + 1) / 200.0
    wh1 = np.ones((1, 2, 8, 8), np.float32)
    out = D.ctdet_decode_batch(*_cuda(heat, wh1), None, down_ratio=1, K=40)
    ref = O.decode(heat, wh1, None, 1, 40)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)


def test_decode_saturated_logits_tie_rule():
    """Logits > 9.21 saturate the clamp at 1-1e-4: probabilities tie, the rule picks lower index."""
    from detectron2_centernet_b200 import decode as D
    hm = np.full((1, 2, 8, 8), -5.0, np.float32)
    hm[0, 1, 3, 3] = 12.0
    hm[0, 0, 5, 5] = 15.0
    heat = S.sigmoid_clamp(hm)
    assert heat[0, 1, 3, 3] == heat[0, 0, 5, 5]
    wh = np.ones((1, 2, 8, 8), np.float32)
    out = D.ctdet_decode_batch(*_cuda(heat, wh), None, down_ratio=1, K=4)
    ref = O.decode(heat, wh, None, 1, 4)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)
    assert out[2][0, 0].item() == 0 and out[3][0, 0].item() == 5 * 8 + 5   # class 0 first (lower flat)


def test_decode_wide_map_strips_and_cat_spec_wh():
    """W = 320 (three 128-column strips incl. a partial one), non-square, cat_spec_wh."""
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 2, 3, 24, 320, 64
    hm, wh, reg = S.make_head_outputs(77, B, C, H, W)
    heat = S.sigmoid_clamp(hm)
    rng = np.random.default_rng(1)
    whc = rng.uniform(0, 32, (B, 2 * C, H, W)).astype(np.float32)
    out = D.ctdet_decode_batch(*_cuda(heat, whc, reg), down_ratio=4, K=K, cat_spec_wh=True)
    ref = O.decode(heat, whc, reg, 4, K, cat_spec_wh=True)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)


def test_decode_odd_width():
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 1, 2, 13, 37, 20
    hm, wh, reg = S.make_head_outputs(78, B, C, H, W)
    heat = S.sigmoid_clamp(hm)
    out = D.ctdet_decode_batch(*_cuda(heat, wh, reg), down_ratio=4, K=K)
    ref = O.decode(heat, wh, reg, 4, K)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)


def test_decode_stream_path_handover_and_workspace_reuse():
    """The streaming decode (seed bound + flat sweep) and its hand-over to the staged kernel in ONE batch:
    image 0/2 = random logits (streaming path), image 1 = what a trained model emits after the clamp of
    centernet.py:161 -- a 1e-4 plateau with a few real peaks, so fewer than K peaks stand above the floor,
    the sampled bound lands on the plateau and the key list overflows -> staged path.  Then the same
    workspace is reused with the roles swapped (it must have been left clean)."""
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 3, 80, 128, 128, 100
    hm, wh, reg = S.make_head_outputs(4321, B, C, H, W)
    heat = S.sigmoid_clamp(hm)
    rng = np.random.default_rng(5)
    plateau = np.full((C, H, W), 1e-4, np.float32)
    for _ in range(30):
        plateau[rng.integers(C), rng.integers(H), rng.integers(W)] = rng.uniform(0.1, 0.9)
    for order in ((0, 1, 0), (1, 0, 1)):
        h = heat.copy()
        for b, flat in enumerate(order):
            if flat:
                h[b] = plateau
        out = D.ctdet_decode_batch(*_cuda(h, wh, reg), down_ratio=4, K=K)
        torch.cuda.synchronize()
        ref = O.decode(h, wh, reg, 4, K)
        for g, r in zip(out, ref):
            assert np.array_equal(g.cpu().numpy(), r)
    # logits mode on the same workspace shape, twice (second call sees whatever the first left behind)
    t = _cuda(hm, wh, reg)
    ref = O.decode(hm, wh, reg, 4, K, logits=True)
    for _ in range(2):
        out = D.ctdet_decode_batch(*t, down_ratio=4, K=K, logits=True)
        for g, r in zip((out[0], out[2], out[3]), (ref[0], ref[2], ref[3])):
            assert np.array_equal(g.cpu().numpy(), r)


def test_decode_stream_path_no_nms_and_small_maps():
    """nms=False (the bare _topk) through the streaming path, and maps too small for a sample (H < 6)."""
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 2, 8, 72, 64, 100
    hm, wh, reg = S.make_head_outputs(99, B, C, H, W)
    out = D.ctdet_decode_batch(*_cuda(hm, wh, reg), down_ratio=4, K=K, logits=True, nms=False)
    flat = torch.from_numpy(hm).view(B, -1)
    top = torch.topk(flat, K)
    assert np.array_equal((out[3].cpu() + out[2].cpu().long() * (H * W)).numpy(), top.indices.numpy())
    B, C, H, W, K = 2, 4, 5, 16, 10
    hm, wh, reg = S.make_head_outputs(98, B, C, H, W)
    heat = S.sigmoid_clamp(hm)
    out = D.ctdet_decode_batch(*_cuda(heat, wh, reg), down_ratio=4, K=K)
    ref = O.decode(heat, wh, reg, 4, K)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)
    # the workspace is shared by every shape with the same B: a staged-only call (above) must leave the regions
    # the streaming layout expects to be zero untouched (regression: its key list used to overlap them)
    B, C, H, W, K = 2, 8, 72, 64, 100
    hm, wh, reg = S.make_head_outputs(97, B, C, H, W)
    out = D.ctdet_decode_batch(*_cuda(hm, wh, reg), down_ratio=4, K=K, logits=True)
    ref = O.decode(hm, wh, reg, 4, K, logits=True)
    assert np.array_equal(out[2].cpu().numpy(), ref[2]) and np.array_equal(out[3].cpu().numpy(), ref[3])
    assert np.array_equal(out[0].cpu().numpy(), ref[0])


def test_decode_concentrated_logits_refined_seed_bound():
    """A random-init head (BASELINE config 5): every logit within a fraction of the -2.19 bias, i.e. the whole
    image inside ONE quarter-octave bin of the bound histogram.  The streaming path refines its sample bound
    below 11 bits instead of handing the image to the staged kernel; results stay exact."""
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 2, 80, 128, 128, 100
    rng = np.random.default_rng(11)
    hm = (-2.19 + 0.1 * rng.standard_normal((B, C, H, W))).astype(np.float32)
    hm[1] = (-2.19 + 1e-3 * rng.standard_normal((C, H, W))).astype(np.float32)     # even tighter
    wh = rng.uniform(0, 64, (B, 2, H, W)).astype(np.float32)
    reg = rng.uniform(0, 1, (B, 2, H, W)).astype(np.float32)
    out = D.ctdet_decode_batch(*_cuda(hm, wh, reg), down_ratio=4, K=K, logits=True)
    ref = O.decode(hm, wh, reg, 4, K, logits=True)
    assert np.array_equal(out[2].cpu().numpy(), ref[2]) and np.array_equal(out[3].cpu().numpy(), ref[3])
    assert np.array_equal(out[0].cpu().numpy(), ref[0])
    heat = S.sigmoid_clamp(hm)
    out = D.ctdet_decode_batch(*_cuda(heat, wh, reg), down_ratio=4, K=K)
    ref = O.decode(heat, wh, reg, 4, K)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)


def test_decode_class_concentrated_scores_use_the_spill_list():
    """Scores concentrated in a few classes (what a trained head emits: a class prior per channel) and around a few
    objects: the candidates above the sampled bound pile into a few tiles, more than a tile's 64-key segment holds.
    Those tiles spill into the image's spill list (one atomic per overflowing tile) instead of handing the whole image to
    the 4x slower staged path; results stay exact.  Image 2 is a plateau: its spill list overflows too -> staged path."""
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 3, 80, 128, 128, 100
    rng = np.random.default_rng(23)
    hm = (0.7 * rng.standard_normal((B, C, H, W)) + (-2.19 + 0.6 * rng.standard_normal((1, C, 1, 1)))).astype(np.float32)
    hm[1, :, 40:56, 30:70] += 2.0                                   # a hot region on top of it
    hm[2] = -2.19
    wh = rng.uniform(0, 64, (B, 2, H, W)).astype(np.float32)
    reg = rng.uniform(0, 1, (B, 2, H, W)).astype(np.float32)
    ref = O.decode(hm, wh, reg, 4, K, logits=True)
    for _ in range(2):                                              # twice: the spill counter must be left at zero
        out = D.ctdet_decode_batch(*_cuda(hm, wh, reg), down_ratio=4, K=K, logits=True)
        assert np.array_equal(out[2].cpu().numpy(), ref[2]) and np.array_equal(out[3].cpu().numpy(), ref[3])
        assert np.array_equal(out[0].cpu().numpy(), ref[0])
    heat = S.sigmoid_clamp(hm)
    out = D.ctdet_decode_batch(*_cuda(heat, wh, reg), down_ratio=4, K=K)
    ref = O.decode(heat, wh, reg, 4, K)
    for g, r in zip(out, ref):
        assert np.array_equal(g.cpu().numpy(), r)


def test_nms_topk_gather_dropins(inputs):
    from detectron2_centernet_b200 import decode as D
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs("dense_small")
    heat = S.sigmoid_clamp(hm)
    (t_heat, t_wh) = _cuda(heat, wh)
    keep = O.peak_mask(heat)
    nms = D._nms(t_heat).cpu().numpy()
    assert np.array_equal(nms, heat * keep)
    s, inds, cls, ys, xs = D._topk(t_heat, K=K)       # bare top-K: no peak test
    flat = heat.reshape(B, -1)
    order = np.argsort(-flat, axis=1, kind="stable")[:, :K]
    assert np.array_equal(s.cpu().numpy(), np.take_along_axis(flat, order, 1))
    assert np.array_equal(cls.cpu().numpy(), (order // (H * W)).astype(np.int32))
    assert np.array_equal(inds.cpu().numpy(), order % (H * W))
    assert np.array_equal(ys.cpu().numpy(), ((order % (H * W)) // W).astype(np.float32))
    g = D._transpose_and_gather_feat(t_wh, inds).cpu().numpy()
    ref = O.gather_feat_nchw(torch.from_numpy(wh), inds.cpu()).numpy()
    assert np.array_equal(g, ref)


def test_topk_dropin_on_arbitrary_scores():
    """`_topk` (centernet.py:408-424) is a bare torch.topk: on inputs that are not post-NMS heat maps -- negative
    scores, zeros, post-NMS maps where most cells are exactly 0 -- it must return the true top K (the
    candidate floor used to drop every score <= 0)."""
    from detectron2_centernet_b200 import decode as D
    rng = np.random.default_rng(5)
    B, C, H, W, K = 2, 6, 40, 64, 50
    for name, x in (("negative", -np.abs(rng.standard_normal((B, C, H, W))).astype(np.float32) - 0.5),
                    ("mixed", rng.standard_normal((B, C, H, W)).astype(np.float32))):
        x = x + (np.arange(C * H * W, dtype=np.float64) / (C * H * W) * 1e-3).astype(np.float32).reshape(C, H, W)   # tie-free
        s, inds, cls, ys, xs = D._topk(torch.from_numpy(x).to(DEV), K=K)
        ref = torch.topk(torch.from_numpy(x).view(B, -1), K)
        assert np.array_equal(s.cpu().numpy(), ref.values.numpy()), name
        assert np.array_equal((cls.cpu().long() * (H * W) + inds.cpu()).numpy(), ref.indices.numpy()), name
    # a post-NMS map: fewer than K positive cells, the rest exactly 0 -- scores must match torch.topk (the order of
    # the tied zeros is unspecified there; ours: lowest flat indices)
    heat = np.zeros((1, 3, 16, 16), np.float32)
    pos = rng.choice(3 * 16 * 16, 20, replace=False)
    heat.reshape(-1)[pos] = rng.uniform(0.1, 1.0, 20).astype(np.float32)
    s, inds, cls, _, _ = D._topk(torch.from_numpy(heat).to(DEV), K=40)
    ref = torch.topk(torch.from_numpy(heat).view(1, -1), 40)
    assert np.array_equal(s.cpu().numpy(), ref.values.numpy())
    flat = (cls.cpu().long() * 256 + inds.cpu()).numpy()[0]
    assert np.array_equal(np.sort(flat[:20]), np.sort(pos)) and len(set(flat.tolist())) == 40


def test_cpu_tensors_raise():
    from detectron2_centernet_b200 import decode as D
    with pytest.raises(NotImplementedError):
        D.ctdet_decode_batch(torch.zeros(1, 2, 8, 8), torch.zeros(1, 2, 8, 8))
    with pytest.raises(RuntimeError):
        D.ctdet_decode_batch(torch.zeros(1, 2, 4, 4, device=DEV), torch.zeros(1, 2, 4, 4, device=DEV), K=17)


# ------------------------------------------------------------------ full-size properties
def test_full_size_properties():
    """BASELINE config 2/3 sizes: size-independent checks (no oracle at this size)."""
    from detectron2_centernet_b200 import decode as D
    from detectron2_centernet_b200 import losses as L
    B, C, H, W, K = 32, 80, 128, 128, 100
    bl, cl = S.make_batch_objects(1236, B, 512)
    gt = _targets(bl, cl, H, W, C)
    hm = gt["hm"]
    assert float(hm.min()) >= 0.0 and float(hm.max()) <= 1.0
    # idempotence / determinism: a second run gives identical bytes
    gt2 = _targets(bl, cl, H, W, C)
    assert torch.equal(hm, gt2["hm"])
    # positives: exactly one 1.0 per distinct (class, centre) -> equals num_pos
    assert int((hm == 1).sum()) == int(gt["num_pos"].sum())
    # image b of the batch equals a batch-of-one run of the same image (image independence)
    one = _targets(bl[5:6], cl[5:6], H, W, C)
    assert torch.equal(one["hm"][0], hm[5]) and torch.equal(one["ind"][0], gt["ind"][5])
    g = torch.Generator(device=DEV).manual_seed(1)
    logits = torch.randn((B, C, H, W), device=DEV, generator=g) * 1.5 - 2.19
    wh = torch.rand((B, 2, H, W), device=DEV, generator=g) * 64
    reg = torch.rand((B, 2, H, W), device=DEV, generator=g)
    # loss linearity in the weights + gradient equals a finite-difference directional derivative
    x = logits.clone().requires_grad_(True)
    o1 = L.ctdet_losses({"hm": x, "wh": wh.clone().requires_grad_(True), "reg": reg.clone().requires_grad_(True)},
                        gt, alpha=[1.0], weights=(1.0, 0.1, 1.0))
    sum(o1.values()).backward()
    o2 = L.ctdet_losses({"hm": logits, "wh": wh, "reg": reg}, gt, alpha=[1.0], weights=(2.0, 0.2, 2.0))
    for k in o1:
        np.testing.assert_allclose(2 * o1[k].item(), o2[k].item(), rtol=1e-6)
    # direction = sign of the gradient: the directional derivative is sum|grad| (large), so the fp32
    # rounding of the two loss values (~2.4e-4 each) stays far below the tolerance and the check is
    # deterministic (a random direction gives a derivative of the size of that noise)
    d = torch.sign(x.grad)
    eps = 1e-2
    lp = L.ctdet_losses({"hm": logits + eps * d, "wh": wh, "reg": reg}, gt, alpha=[1.0])["hm_loss"].item()
    lm = L.ctdet_losses({"hm": logits - eps * d, "wh": wh, "reg": reg}, gt, alpha=[1.0])["hm_loss"].item()
    fd = (lp - lm) / (2 * eps)
    an = float((x.grad.double() * d.double()).sum())
    np.testing.assert_allclose(an, fd, rtol=5e-3)   # central difference: truncation ~eps^2
    # decode: scores sorted, every winner is a 3x3 peak, nothing outside the winners beats the K-th
    heat = torch.clamp(torch.sigmoid(logits), 1e-4, 1 - 1e-4)
    boxes, scores, classes, inds = D.ctdet_decode_batch(heat, wh, reg, down_ratio=4, K=K)
    assert bool((scores[:, :-1] >= scores[:, 1:]).all())
    pooled = torch.nn.functional.max_pool2d(heat, 3, 1, 1)
    peaks = torch.where(pooled == heat, heat, torch.zeros_like(heat)).view(B, -1)
    flat = classes.long() * (H * W) + inds
    assert torch.equal(peaks.gather(1, flat), scores)
    ref_s, ref_i = torch.topk(peaks, K, dim=1)                 # tie-free random data
    assert torch.equal(ref_s, scores) and torch.equal(ref_i, flat)
    # batch independence: image 7 alone decodes to the same detections
    b1, s1, c1, i1 = D.ctdet_decode_batch(heat[7:8], wh[7:8], reg[7:8], down_ratio=4, K=K)
    assert torch.equal(b1[0], boxes[7]) and torch.equal(c1[0], classes[7])


def test_config4_1024_targets_and_decode():
    """BASELINE config 4: B = 8 at 1024^2 -> [8, 80, 256, 256] maps, 128 objects per image (the MAXOBJ
    limit of detection_utils.py:606).  Two images are checked against the oracle outright (the oracle
    needs ~50 ms per image at this size), the whole batch through size-independent properties."""
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 8, 80, 256, 256, 100
    bl, cl = S.make_batch_objects(1238, B, 1024, n=128)
    out = _targets(bl, cl, H, W, C)
    ref = O.gen_heatmap_batch(bl[:2], cl[:2], H, W, C)
    for k in ("reg_mask", "ind", "wh", "reg"):
        assert np.array_equal(out[k][:2].cpu().numpy(), ref[k]), k
    got = out["hm"][:2].cpu().numpy()
    assert np.abs(got - ref["hm"]).max() <= 1e-6
    assert np.array_equal(got == 1, ref["hm"] == 1) and np.array_equal(got == 0, ref["hm"] == 0)
    assert int((out["hm"] == 1).sum()) == int(out["num_pos"].sum())
    assert int(out["reg_mask"].sum()) == sum(int(((b[:, 2] > b[:, 0]) & (b[:, 3] > b[:, 1])).sum()) for b in bl)
    hm, wh, reg = S.make_head_outputs(1238, B, C, H, W, boxes_list=bl)
    heat = S.sigmoid_clamp(hm)
    heat_t, wh_t, reg_t = _cuda(heat, wh, reg)
    boxes, scores, classes, inds = D.ctdet_decode_batch(heat_t, wh_t, reg_t, down_ratio=4, K=K)
    for b in range(2):                                              # oracle, image by image like the reference
        rb, rs, rc, ri = O.decode(heat[b:b + 1], wh[b:b + 1], reg[b:b + 1], 4, K)
        assert np.array_equal(scores[b].cpu().numpy(), rs[0]) and np.array_equal(classes[b].cpu().numpy(), rc[0])
        assert np.array_equal(inds[b].cpu().numpy(), ri[0]) and np.array_equal(boxes[b].cpu().numpy(), rb[0])
    pooled = torch.nn.functional.max_pool2d(heat_t, 3, 1, 1)
    peaks = torch.where(pooled == heat_t, heat_t, torch.zeros_like(heat_t)).view(B, -1)
    ref_s, ref_i = torch.topk(peaks, K, dim=1)
    assert torch.equal(ref_s, scores) and torch.equal(ref_i, classes.long() * (H * W) + inds)


def test_config3_b64_decode():
    """BASELINE config 3: B = 64 decode, K = 100, 3x3 NMS over [64, 80, 128, 128]."""
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 64, 80, 128, 128, 100
    g = torch.Generator(device=DEV).manual_seed(3)
    logits = torch.randn((B, C, H, W), device=DEV, generator=g) * 1.5 - 2.19
    wh = torch.rand((B, 2, H, W), device=DEV, generator=g) * 64
    reg = torch.rand((B, 2, H, W), device=DEV, generator=g)
    heat = torch.clamp(torch.sigmoid(logits), 1e-4, 1 - 1e-4)
    boxes, scores, classes, inds = D.ctdet_decode_batch(heat, wh, reg, down_ratio=4, K=K)
    pooled = torch.nn.functional.max_pool2d(heat, 3, 1, 1)
    peaks = torch.where(pooled == heat, heat, torch.zeros_like(heat)).view(B, -1)
    ref_s, ref_i = torch.topk(peaks, K, dim=1)
    assert torch.equal(ref_s, scores) and torch.equal(ref_i, classes.long() * (H * W) + inds)
    # boxes from the gathered wh / reg, in the reference's operation order (centernet.py:436-455)
    xs = (inds % W).float() + reg[:, 0].reshape(B, -1).gather(1, inds)
    ys = (inds // W).float() + reg[:, 1].reshape(B, -1).gather(1, inds)
    w2 = wh[:, 0].reshape(B, -1).gather(1, inds) / 2
    h2 = wh[:, 1].reshape(B, -1).gather(1, inds) / 2
    want = torch.stack([xs - w2, ys - h2, xs + w2, ys + h2], dim=2) * 4
    assert torch.equal(boxes, want)
    # one image through the oracle
    rb, rs, rc, ri = O.decode(heat[63:64].cpu().numpy(), wh[63:64].cpu().numpy(), reg[63:64].cpu().numpy(), 4, K)
    assert np.array_equal(scores[63].cpu().numpy(), rs[0]) and np.array_equal(classes[63].cpu().numpy(), rc[0])
    assert np.array_equal(boxes[63].cpu().numpy(), rb[0])


def test_postprocess_batch():
    from detectron2_centernet_b200 import decode as D
    B, C, H, W, K = 3, 5, 32, 32, 40
    hm, wh, reg = S.make_head_outputs(5, B, C, H, W)
    heat = S.sigmoid_clamp(hm)
    boxes, scores, classes, inds = D.ctdet_decode_batch(*_cuda(heat, wh, reg), down_ratio=4, K=K)
    rb, rs, rc = boxes.cpu().numpy().copy(), scores.cpu().numpy().copy(), classes.cpu().numpy().copy()
    sizes = np.array([[128, 128, 256, 192], [128, 128, 128, 128], [100, 120, 50, 60]], np.int32)
    n_keep = D.postprocess_batch(boxes, scores, classes, torch.from_numpy(sizes).to(DEV),
                                 max_dets=30, score_thresh=0.3)
    for b in range(B):
        bb, ss, cc = O.postprocess_single(rb[b], rs[b], rc[b], 30, 0.3)
        in_h, in_w, out_h, out_w = [int(v) for v in sizes[b]]
        t = torch.from_numpy(bb.copy())
        t[:, 0::2] *= out_w / in_w
        t[:, 1::2] *= out_h / in_h
        t[:, 0::2].clamp_(0, out_w)
        t[:, 1::2].clamp_(0, out_h)
        ne = ((t[:, 2] - t[:, 0]) > 0) & ((t[:, 3] - t[:, 1]) > 0)
        n = int(ne.sum())
        assert int(n_keep[b]) == n
        assert np.array_equal(boxes[b, :n].cpu().numpy(), t[ne].numpy())
        assert np.array_equal(scores[b, :n].cpu().numpy(), ss[ne.numpy()])
        assert np.array_equal(classes[b, :n].cpu().numpy(), cc[ne.numpy()])


# ------------------------------------------------------------------ meta-arch glue
def test_centernet_hot_path_losses_and_inference(inputs):
    """CenterNetHotPath.losses / inference against the oracle's restatement of
    CenterNet.losses (:191-212) and inference_single_image (:236-266) + detector_postprocess."""
    import detectron2_centernet_b200 as P
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs("coco512")
    model = P.CenterNetHotPath(P.get_cfg(overrides={"MODEL": {"CENTERNET": {"FOCAL_LOSS_ALPHA": [1]}}}), num_classes=C)
    insts = []
    for b in range(B):
        i = P.Instances((512, 512))
        i.gt_boxes = P.Boxes(torch.from_numpy(bl[b]).to(DEV))
        i.gt_classes = torch.from_numpy(cl[b]).to(DEV)
        insts.append(i)
    targets = model.targets(insts, (512, 512))
    t_hm, t_wh, t_reg = [t.requires_grad_(True) for t in _cuda(hm, wh, reg)]
    out = model.losses({"hm": t_hm, "wh": t_wh, "reg": t_reg}, targets)
    gt_np = O.gen_heatmap_batch(bl, cl, H, W, C)
    ref, _ = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), gt_np, alpha=[1.0])
    for k in ("hm_loss", "wh_loss", "off_loss"):
        np.testing.assert_allclose(out[k].item(), float(ref[k]), rtol=1e-5)
    # the optional all-reduce path with one process gives the same numbers
    model_sync = P.CenterNetHotPath(P.get_cfg(overrides={"MODEL": {"CENTERNET": {"FOCAL_LOSS_ALPHA": [1]}}}),
                                    num_classes=C, sync_num_pos=True)
    out2 = model_sync.losses({"hm": t_hm, "wh": t_wh, "reg": t_reg}, targets)
    for k in out:
        np.testing.assert_allclose(out2[k].item(), out[k].item(), rtol=1e-6)
    # one-pass variant: targets generated inside the loss sweep (no heatmap target in HBM)
    out3, slots3 = model.losses_from_instances({"hm": t_hm, "wh": t_wh, "reg": t_reg}, insts)
    out4, _ = model_sync.losses_from_instances({"hm": t_hm, "wh": t_wh, "reg": t_reg}, insts)
    for k in out:
        np.testing.assert_allclose(out3[k].item(), float(ref[k]), rtol=1e-5)
        np.testing.assert_allclose(out4[k].item(), float(ref[k]), rtol=1e-5)
    assert np.array_equal(slots3["ind"].cpu().numpy(), gt_np["ind"])
    # inference: list of Instances, thresholded (score > 0.05) and resized to the original image size
    sizes, orig = [(512, 512)] * B, [(1024, 768), (512, 512)][:B]
    res = model.inference({"hm": t_hm.detach(), "wh": t_wh.detach(), "reg": t_reg.detach()}, sizes, orig)
    ob, os_, oc, _ = O.decode(S.sigmoid_clamp(hm), wh, reg, 4, model.topk_candidates)
    for b in range(B):
        bb, ss, cc = O.postprocess_single(ob[b], os_[b], oc[b], model.max_detections_per_image, model.score_threshold)
        t = torch.from_numpy(bb.copy())
        t[:, 0::2] *= orig[b][1] / sizes[b][1]
        t[:, 1::2] *= orig[b][0] / sizes[b][0]
        t[:, 0::2].clamp_(0, orig[b][1])
        t[:, 1::2].clamp_(0, orig[b][0])
        keep = ((t[:, 2] - t[:, 0]) > 0) & ((t[:, 3] - t[:, 1]) > 0)
        r = res[b]
        assert r.image_size == orig[b] and len(r) == int(keep.sum())
        assert r.pred_classes.dtype == torch.int32
        assert np.array_equal(r.pred_classes.cpu().numpy(), cc[keep.numpy()])
        np.testing.assert_allclose(r.scores.cpu().numpy(), ss[keep.numpy()], rtol=2e-7)
        assert np.array_equal(r.pred_boxes.tensor.cpu().numpy(), t[keep].numpy())


def test_fused_losses_second_backward_raises_clearly(inputs):
    """The three losses share one autograd node whose gradients are released by the first backward."""
    import detectron2_centernet_b200 as P
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs("tiny")
    boxes, classes, offsets = S.pack_objects(bl, cl)
    gt = P.target.target_gen(*_cuda(boxes, classes, offsets), B, C, H, W)
    t = [a.requires_grad_(True) for a in _cuda(hm, wh, reg)]
    out = P.losses.ctdet_losses({"hm": t[0], "wh": t[1], "reg": t[2]}, gt, alpha=[1.0])
    out["hm_loss"].backward(retain_graph=True)
    with pytest.raises(RuntimeError, match="already back-propagated"):
        out["wh_loss"].backward()


def test_config2_full_size_against_the_oracle():
    """BASELINE config 2 at its real size -- B = 32 @512^2, the bench's own inputs (seeds 1236 / 2236) -- compared with
    the pinned oracle, not only through properties: targets (hm <= 1e-6, slots bit-exact), the three
    losses (1e-5 relative) and the dense hm gradient (1e-5 relative to its scale), for the two-kernel form AND the
    fused target+loss sweep; and the decode of the same batch in logits mode (classes / indices / boxes bit-exact)."""
    import detectron2_centernet_b200 as P
    B, C, H, W, K = 32, 80, 128, 128, 100
    bl, cl = S.make_batch_objects(1236, B, 512)
    hm, wh, reg = S.make_head_outputs(2236, B, C, H, W, bl, tie_free=True)
    boxes, classes, offsets = S.pack_objects(bl, cl)
    t_b, t_c, t_o = _cuda(boxes, classes, offsets)
    gt = P.target.target_gen(t_b, t_c, t_o, B, C, H, W)
    ref_gt = O.gen_heatmap_batch(bl, cl, H, W, C)
    assert float(np.abs(gt["hm"].cpu().numpy() - ref_gt["hm"]).max()) <= 1e-6
    for k in ("ind", "reg_mask", "wh", "reg"):
        assert np.array_equal(gt[k].cpu().numpy(), ref_gt[k]), k
    torch.set_num_threads(max(1, torch.get_num_threads()))
    ref_l, ref_g = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), ref_gt, alpha=[1.0])
    scale = float(ref_g[0].abs().max())
    for fused in (False, True):
        t = [a.requires_grad_(True) for a in _cuda(hm, wh, reg)]
        if fused:
            out, _ = P.losses.ctdet_losses_from_objects({"hm": t[0], "wh": t[1], "reg": t[2]}, t_b, t_c, t_o, alpha=[1.0])
        else:
            out = P.losses.ctdet_losses({"hm": t[0], "wh": t[1], "reg": t[2]}, gt, alpha=[1.0])
        sum(out.values()).backward()
        for k in ref_l:
            np.testing.assert_allclose(out[k].item(), float(ref_l[k]), rtol=1e-5, err_msg="%s fused=%s" % (k, fused))
        err = float((t[0].grad.cpu() - ref_g[0]).abs().max())
        assert err <= 1e-5 * scale, (fused, err, scale)
        for g, r in ((t[1].grad, ref_g[1]), (t[2].grad, ref_g[2])):
            assert float((g.cpu() - r).abs().max()) <= 1e-5 * max(float(r.abs().max()), 1e-12)
    det = P.decode.ctdet_decode_batch(*_cuda(hm, wh, reg), down_ratio=4, K=K, logits=True)
    rb, rs, rc, ri = O.decode(hm, wh, reg, 4, K, logits=True)
    assert np.array_equal(det[2].cpu().numpy(), rc) and np.array_equal(det[3].cpu().numpy(), ri)
    assert np.array_equal(det[0].cpu().numpy(), rb)
    np.testing.assert_allclose(det[1].cpu().numpy(), rs, rtol=2e-7)


def test_tma_staged_loss_sweep_in_a_fresh_process():
    """`CTDET_LOSS_TMA=1` (read once per process) selects loss_kernel_tma -- the sweep's reads staged by bulk async copies
    into a shared-memory ring (opt-in: measured slower than the register-pipelined sweep) -- with a
    2-stage ring of 4 float4 per thread, so chunks end inside planes and the last chunk is partial: the same loss parity
    tests in a child process."""
    import os
    import subprocess
    import sys
    if os.environ.get("CTDET_LOSS_TMA"):
        pytest.skip("already the child")
    env = dict(os.environ, CTDET_LOSS_TMA="1", CTDET_LOSS_TMA_Q="4", CTDET_LOSS_TMA_STAGES="2")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu", "-k",
                        "test_fused_losses or edge_shapes or no_positives or dropins"], env=env, capture_output=True,
                       text=True, timeout=300, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
