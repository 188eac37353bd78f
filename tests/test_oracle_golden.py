"""CPU: the oracle (oracle/ctdet_oracle.py) against the golden vectors produced by the reference's
own functions (oracle/gen_golden.py), and against the live reference when its tree is present."""
import hashlib

import numpy as np
import pytest
import torch

from oracle import ctdet_oracle as O
from oracle import gen_golden as G
from oracle import ref_loader as RL
from detectron2_centernet_b200 import synthetic as S

CASES = list(G.CASES)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def inputs():
    cache = {}

    def get(case):
        if case not in cache:
            cache[case] = G.build_inputs(case)
        return cache[case]
    return get


@pytest.mark.parametrize("case", CASES)
def test_inputs_reproduce(case, golden, inputs):
    bl, cl, hm, wh, reg, _ = inputs(case)
    boxes, classes, _ = S.pack_objects(bl, cl)
    got = [sha(boxes), sha(classes), sha(hm), sha(wh), sha(reg)]
    assert got == list(golden[case + "/in_sha"]), "synthetic generator drifted"


@pytest.mark.parametrize("case", CASES)
def test_target_gen_matches_reference(case, golden, inputs):
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs(case)
    gt = O.gen_heatmap_batch(bl, cl, H, W, C)
    for k in ("reg_mask", "ind", "wh", "reg"):
        ref = golden[case + "/gt_" + k]
        assert gt[k].dtype == ref.dtype and np.array_equal(gt[k], ref), k
    assert sha(gt["hm"]) == str(golden[case + "/gt_hm_sha"])          # bit-exact heatmap
    assert O.count_positives(gt) == int(golden[case + "/num_pos"])
    if full:
        assert np.array_equal(gt["hm"], golden[case + "/gt_hm"])


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("tag", ["a1", "acls"])
def test_losses_match_reference(case, tag, golden, inputs):
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs(case)
    alpha = [1.0] if tag == "a1" else [0.25 + 0.01 * c for c in range(C)]
    gt = O.gen_heatmap_batch(bl, cl, H, W, C)
    torch.set_num_threads(1)
    lo, gr = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg),
                                gt, alpha=alpha)
    got = np.array([float(lo["hm_loss"]), float(lo["wh_loss"]), float(lo["off_loss"])])
    np.testing.assert_allclose(got, golden["%s/loss_%s" % (case, tag)], rtol=1e-5)   # north_star: 1e-5 rel
    for nm, g in zip(("dhm", "dwh", "dreg"), gr):
        g = g.numpy()
        if full:
            ref = golden["%s/%s_%s" % (case, nm, tag)]
            np.testing.assert_allclose(g, ref, rtol=1e-5, atol=1e-6 * np.abs(ref).max())
        else:
            ref = golden["%s/%s_%s_sample" % (case, nm, tag)]
            np.testing.assert_allclose(g.reshape(-1)[::997], ref, rtol=1e-5, atol=1e-6 * np.abs(ref).max())
            np.testing.assert_allclose(np.abs(g.astype(np.float64)).sum(),
                                       float(golden["%s/%s_%s_l1" % (case, nm, tag)]), rtol=1e-5)


@pytest.mark.parametrize("case", CASES)
def test_decode_matches_reference(case, golden, inputs):
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = inputs(case)
    heat = S.sigmoid_clamp(hm)
    assert sha(heat) == str(golden[case + "/heat_sha"])
    boxes, scores, classes, inds = O.decode(heat, wh, reg, down_ratio=4, K=K)
    assert np.array_equal(classes, golden[case + "/dec_classes"]) and classes.dtype == np.int32
    assert np.array_equal(scores, golden[case + "/dec_scores"])
    assert np.array_equal(boxes, golden[case + "/dec_boxes"])            # bit-exact boxes
    b2, _, _, _ = O.decode(heat, wh, None, down_ratio=4, K=K)
    assert np.array_equal(b2, golden[case + "/dec_boxes_noreg"])
    bt, st, ct, it = O.decode_torch(heat, wh, reg, down_ratio=4, K=K)      # the timed CPU-baseline variant
    assert np.array_equal(bt, boxes) and np.array_equal(st, scores) and np.array_equal(ct, classes)
    # ranking on raw logits is the same ranking (sigmoid is monotone, inputs are tie-free)
    b3, s3, c3, i3 = O.decode(hm, wh, reg, down_ratio=4, K=K, logits=True)
    assert np.array_equal(c3, classes) and np.array_equal(i3, inds) and np.array_equal(b3, boxes)
    np.testing.assert_allclose(s3, scores, rtol=2e-7)


@pytest.mark.skipif(not RL.available(), reason="reference tree not present (GPU box)")
def test_oracle_vs_live_reference_random():
    """Fresh seeds not in the golden file, straight against the reference's functions."""
    for seed in (11, 12, 13):
        B, size, C = 2, 256, 7
        H = W = size // 4
        bl, cl = S.make_batch_objects(seed, B, size, n_max=60, num_classes=C)
        gt = O.gen_heatmap_batch(bl, cl, H, W, C)
        for b in range(B):
            r = RL.ref_gen_heatmap(bl[b], cl[b], H, W, C)
            for k in r:
                assert np.array_equal(r[k].numpy(), gt[k][b]), (seed, b, k)
        hm, wh, reg = S.make_head_outputs(seed, B, C, H, W, bl)
        heat = S.sigmoid_clamp(hm)
        rb, rs, rc = RL.ref_decode(torch.from_numpy(heat), torch.from_numpy(wh), torch.from_numpy(reg), 4, 50)
        ob, os_, oc, _ = O.decode(heat, wh, reg, 4, 50)
        assert np.array_equal(rb.numpy(), ob) and np.array_equal(rs.numpy(), os_) and np.array_equal(rc.numpy(), oc)


def test_gaussian_radius_known_values():
    # hand-evaluated from the three quadratics of detection_utils.py:654-680
    assert int(O.gaussian_radius(1, 1)) == 0
    assert int(O.gaussian_radius(10, 10)) == 2
    assert int(O.gaussian_radius(50, 30)) == 10    # (-112 + sqrt(12544 + 5040)) / 2 = 10.30
    r = O.gaussian_radius(64, 64)
    assert abs(r - min((128 + np.sqrt(128 ** 2 - 4 * 64 * 64 * 0.3 / 1.7)) / 2,
                       (256 + np.sqrt(256 ** 2 - 16 * 0.3 * 64 * 64)) / 2,
                       (-1.4 * 128 + np.sqrt((1.4 * 128) ** 2 + 4 * 2.8 * 0.3 * 64 * 64)) / 2)) < 1e-9


def test_decode_fewer_peaks_than_k():
    """heat*keep zeros fill the tail (centernet.py:405): by our tie rule the lowest flat indices."""
    heat = np.full((1, 2, 4, 4), 0.1, np.float32)
    heat[0, 1, 2, 2] = 0.9
    heat[0, 0, 0, 0] = 0.5
    wh = np.ones((1, 2, 4, 4), np.float32)
    boxes, scores, classes, inds = O.decode(heat, wh, None, 1, K=5)
    assert scores[0, 0] == np.float32(0.9) and classes[0, 0] == 1 and inds[0, 0] == 10
    assert scores[0, 1] == np.float32(0.5) and classes[0, 1] == 0 and inds[0, 1] == 0
    # plateau of 0.1 cells: all of them are peaks unless next to a larger value
    assert np.all(scores[0, 2:] == np.float32(0.1))
