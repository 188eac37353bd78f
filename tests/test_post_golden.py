"""Inference glue pinned on the reference: golden vectors in
tests/golden/postprocess_golden.npz were produced by the reference's OWN `CenterNet.inference` /
`inference_single_image` (centernet.py:214-266), the eval-branch sigmoid_ + clamp (:161) and
`detector_postprocess` (modeling/postprocessing.py:11-74) with the reference's own Instances / Boxes
(oracle/gen_golden_post.py).

  CPU   the oracle restatement (oracle/ctdet_oracle.py: decode, postprocess_single, detector_postprocess,
        inference) against the golden vectors, and the live reference when its tree is present
  GPU   `ctdet_postprocess` / `CenterNetHotPath.inference` (logits mode: the product's hot path) against the
        same vectors

Tolerances: counts, classes and the order of detections exact on the tie-free cases; boxes bit-exact; scores
within 2e-7 relative (the reference's CPU sigmoid vs `1/(1+expf(-x))`).  The `plateau` case has tied scores by
construction (the reference's torch.topk order among ties is unspecified): compared as sorted sets.
"""
import os

import numpy as np
import pytest
import torch

from oracle import ctdet_oracle as O
from oracle import gen_golden_post as GP
from oracle import ref_loader as RL

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = "cuda:0"


@pytest.fixture(scope="module")
def pg():
    return np.load(os.path.join(ROOT, "tests", "golden", "postprocess_golden.npz"))


def _split(pg, name, stage):
    counts = pg["%s/%s/counts" % (name, stage)]
    ends = np.cumsum(counts)
    starts = ends - counts
    out = []
    for s, e in zip(starts, ends):
        out.append((pg["%s/%s/boxes" % (name, stage)][s:e], pg["%s/%s/scores" % (name, stage)][s:e],
                    pg["%s/%s/classes" % (name, stage)][s:e]))
    return out


def _canon(boxes, scores, classes):
    """order-independent form for the tied case: rows sorted by (class, box)"""
    rows = np.concatenate([np.asarray(classes, np.float64)[:, None], np.asarray(boxes, np.float64),
                           np.asarray(scores, np.float64)[:, None]], axis=1)
    return rows[np.lexsort(rows.T[::-1])]


def _compare(name, got, want):
    assert len(got) == len(want)
    for i, ((gb, gs, gc), (wb, ws, wc)) in enumerate(zip(got, want)):
        assert len(gs) == len(ws), (name, i, len(gs), len(ws))
        if name == "plateau":
            g, w = _canon(gb, gs, gc), _canon(wb, ws, wc)
            assert np.array_equal(g[:, :5], w[:, :5]), (name, i)
            np.testing.assert_allclose(g[:, 5], w[:, 5], rtol=2e-7)
            np.testing.assert_allclose(np.sort(gs)[::-1], gs, rtol=0)        # still sorted by score
        else:
            assert np.array_equal(gc, wc), (name, i)
            assert np.asarray(gc).dtype == np.int32
            assert np.array_equal(gb, wb), (name, i)
            np.testing.assert_allclose(gs, ws, rtol=2e-7)


# ------------------------------------------------------------------------------------------------ CPU
@pytest.mark.parametrize("name", GP.CASES)
@pytest.mark.parametrize("stage", ["raw", "post"])
def test_oracle_inference_matches_reference_golden(pg, name, stage):
    c = GP.build_case(name)
    got = O.inference(c["hm"], c["wh"], c["reg"], c["image_sizes"], c["output_sizes"] if stage == "post" else None,
                      **c["cfg"])
    _compare(name, got, _split(pg, name, stage))


@pytest.mark.parametrize("name", GP.CASES)
def test_oracle_logits_mode_equals_probability_mode(name):
    """decode(logits=True) -- the rule the kernel implements -- yields the detections of the reference chain
    (sigmoid + clamp, then decode on probabilities), plateaus and clamp saturation included."""
    c = GP.build_case(name)
    K = c["cfg"]["topk_candidates"]
    b1, s1, c1, _ = O.decode(c["hm"], c["wh"], c["reg"], 4, K, logits=True)
    b2, s2, c2, _ = O.decode(O._sigmoid_clamp(c["hm"]), c["wh"], c["reg"], 4, K)
    for i in range(len(s1)):
        k1, k2 = s1[i] > 0.05, s2[i] > 0.05
        assert np.array_equal(_canon(b1[i][k1], s1[i][k1], c1[i][k1]), _canon(b2[i][k2], s2[i][k2], c2[i][k2]))


@pytest.mark.skipif(not RL.available(), reason="reference tree not present")
@pytest.mark.parametrize("name", GP.CASES)
def test_live_reference_matches_golden(pg, name):
    c = GP.build_case(name)
    t = [torch.from_numpy(c[k]) for k in ("hm", "wh", "reg")]
    for stage, osz in (("raw", None), ("post", c["output_sizes"])):
        res = RL.ref_inference(t[0], t[1], t[2], c["image_sizes"], osz, **c["cfg"])
        got = [(r.pred_boxes.tensor.numpy(), r.scores.numpy(), r.pred_classes.numpy()) for r in res]
        want = _split(pg, name, stage)
        for (gb, gs, gc), (wb, ws, wc) in zip(got, want):
            assert np.array_equal(gb, wb) and np.array_equal(gs, ws) and np.array_equal(gc, wc)
        sizes = [tuple(int(v) for v in r.image_size) for r in res]
        assert sizes == [tuple(int(v) for v in s) for s in pg["%s/%s/image_sizes" % (name, stage)]]


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", GP.CASES)
@pytest.mark.parametrize("stage", ["raw", "post"])
def test_hot_path_inference_matches_reference_golden(pg, name, stage):
    import detectron2_centernet_b200 as P
    c = GP.build_case(name)
    cfg = c["cfg"]
    model = P.CenterNetHotPath(P.get_cfg(overrides={
        "MODEL": {"CENTERNET": {"SCORE_THRESH_TEST": cfg["score_threshold"], "TOPK_CANDIDATES_TEST": cfg["topk_candidates"],
                                "DOWN_RATIO": cfg["down_ratio"]}},
        "TEST": {"DETECTIONS_PER_IMAGE": cfg["max_detections_per_image"]}}), num_classes=c["hm"].shape[1])
    out = {k: torch.from_numpy(c[k]).to(DEV) for k in ("hm", "wh", "reg")}
    res = model.inference(out, c["image_sizes"], c["output_sizes"] if stage == "post" else None)
    got = [(r.pred_boxes.tensor.cpu().numpy(), r.scores.cpu().numpy(), r.pred_classes.cpu().numpy()) for r in res]
    _compare(name, got, _split(pg, name, stage))
    want_sizes = [tuple(int(v) for v in s) for s in pg["%s/%s/image_sizes" % (name, stage)]]
    assert [tuple(r.image_size) for r in res] == want_sizes
    for r in res:
        assert r.pred_classes.dtype == torch.int32 and isinstance(r.pred_boxes, P.Boxes)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["mixed", "coco512"])
def test_postprocess_kernel_on_probability_decode_matches_reference_golden(pg, name):
    """The drop-in chain a reference user runs: heat = clamp(sigmoid) -> ctdet_decode_batch (probability mode) ->
    ctdet_postprocess."""
    from detectron2_centernet_b200 import decode as D
    from detectron2_centernet_b200 import synthetic as S
    c = GP.build_case(name)
    cfg = c["cfg"]
    heat = torch.from_numpy(S.sigmoid_clamp(c["hm"])).to(DEV)
    boxes, scores, classes, _ = D.ctdet_decode_batch(heat, torch.from_numpy(c["wh"]).to(DEV), torch.from_numpy(c["reg"]).to(DEV),
                                                     down_ratio=cfg["down_ratio"], K=cfg["topk_candidates"])
    sizes = torch.tensor([[i[0], i[1], o[0], o[1]] for i, o in zip(c["image_sizes"], c["output_sizes"])],
                         dtype=torch.int32, device=DEV)
    n_keep = D.postprocess_batch(boxes, scores, classes, sizes, cfg["max_detections_per_image"],
                                 cfg["score_threshold"]).cpu().numpy()
    got = [(boxes[b, :n].cpu().numpy(), scores[b, :n].cpu().numpy(), classes[b, :n].cpu().numpy())
           for b, n in enumerate(n_keep)]
    _compare(name, got, _split(pg, name, "post"))
