"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/postprocess_golden.npz by running the reference's OWN,
unmodified `CenterNet.inference` / `inference_single_image` (meta_arch/centernet.py:214-266), the eval-branch
sigmoid_ + clamp (:161) and `detector_postprocess` (modeling/postprocessing.py:11-74) with the reference's own
`Instances` / `Boxes` (oracle/ref_loader.py) on seeded inputs.

Run in the build container (needs /root/reference):   python -m oracle.gen_golden_post

Cases (inputs are rebuilt from the seeds by `build_case`, which the tests import; outputs stored in full):
  mixed        B=5, C=6, 32x32, K=20: plain image; non-square up-scale; down-scale with boxes clipped away;
               image whose scores all sit under the threshold (empty result); max_detections_per_image < K
  coco512      B=2, 80x128x128, K=100, rescaled to (1024, 768) and (333, 500)
  plateau      logits mode, crafted: saturated neighbours (10 next to 11: both clamp to 1-1e-4, the
               reference keeps BOTH), near-equal neighbours below the clamp (x and nextafter(x): same fp32
               probability), exactly equal neighbours, a saturated 2x2 block
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_loader as RL                               # noqa: E402
from detectron2_centernet_b200 import synthetic as S              # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "tests", "golden", "postprocess_golden.npz")


def robust_collision_logit(lo, hi):
    """A logit x in (lo, hi) such that sigmoid(x) and sigmoid(nextafter(x)) round to the SAME fp32 value under
    any sigmoid implementation accurate to 0.25 ulp of the result: both exact values sit in the middle half of
    one fp32 rounding interval."""
    x = np.float32(lo)
    while x < hi:
        xn = np.nextafter(x, np.float32(np.inf))
        ok = True
        cells = []
        for v in (x, xn):
            p = 1.0 / (1.0 + np.exp(-np.float64(v)))
            pf = np.float32(p)
            ulp = np.float64(np.spacing(pf))
            frac = (p - np.float64(pf)) / ulp            # in [-0.5, 0.5]
            ok &= abs(frac) < 0.25
            cells.append(pf)
        if ok and cells[0] == cells[1]:
            return x, xn
        x = np.nextafter(xn, np.float32(np.inf))
    raise RuntimeError("no robust collision found")


def build_case(name):
    """-> dict(hm logits, wh, reg, image_sizes, output_sizes, cfg)"""
    if name == "mixed":
        B, C, H, W = 5, 6, 32, 32
        hm, wh, reg = S.make_head_outputs(77, B, C, H, W, tie_free=True)
        hm = hm.copy()
        hm[3] -= 9.0                                   # every probability < 0.05: empty result
        wh = wh.copy()
        wh[2, :, :, :16] = 0.0                         # zero-size boxes in the left half: dropped by nonempty()
        return dict(hm=hm, wh=wh, reg=reg,
                    image_sizes=[(128, 128), (128, 128), (128, 128), (128, 128), (100, 120)],
                    output_sizes=[(128, 128), (300, 200), (20, 24), (128, 128), (50, 60)],
                    cfg=dict(down_ratio=4, topk_candidates=20, max_detections_per_image=12, score_threshold=0.05))
    if name == "coco512":
        B, C, H, W = 2, 80, 128, 128
        bl, _ = S.make_batch_objects(1235, B, 512)
        hm, wh, reg = S.make_head_outputs(1242, B, C, H, W, bl, tie_free=True)
        return dict(hm=hm, wh=wh, reg=reg, image_sizes=[(512, 512), (512, 512)],
                    output_sizes=[(1024, 768), (333, 500)],
                    cfg=dict(down_ratio=4, topk_candidates=100, max_detections_per_image=100, score_threshold=0.05))
    if name == "plateau":
        B, C, H, W = 2, 3, 16, 16
        rng = np.random.default_rng(9)
        hm = rng.normal(-6.0, 0.5, (B, C, H, W)).astype(np.float32)       # background: p ~ 0.0025 < threshold
        wh = rng.uniform(2.0, 12.0, (B, 2, H, W)).astype(np.float32)
        reg = rng.uniform(0.0, 1.0, (B, 2, H, W)).astype(np.float32)
        a, an = robust_collision_logit(5.5, 6.5)
        # image 0
        hm[0, 0, 3, 3], hm[0, 0, 3, 4] = 10.0, 11.0                       # both saturate: two detections
        hm[0, 1, 8, 8], hm[0, 1, 9, 9] = a, an                            # same fp32 probability below the clamp
        hm[0, 2, 5, 5], hm[0, 2, 5, 6] = 2.0, 2.0                         # exactly equal neighbours
        hm[0, 2, 12, 3], hm[0, 2, 12, 4] = 3.0, 2.5                       # ordinary suppression: one detection
        # image 1: a saturated 2x2 block and a lone strong peak next to a slightly weaker saturated one
        hm[1, 0, 6:8, 6:8] = np.array([[9.5, 12.0], [15.0, 9.3]], np.float32)
        hm[1, 1, 2, 2], hm[1, 1, 2, 3] = 20.0, 9.25
        hm[1, 2, 10, 10] = 1.0
        return dict(hm=hm, wh=wh, reg=reg, image_sizes=[(64, 64), (64, 64)], output_sizes=[(64, 64), (128, 96)],
                    cfg=dict(down_ratio=4, topk_candidates=40, max_detections_per_image=40, score_threshold=0.05))
    raise KeyError(name)


CASES = ("mixed", "coco512", "plateau")


def main():
    assert RL.available(), "needs the reference tree"
    torch.set_num_threads(1)
    out = {}
    for name in CASES:
        c = build_case(name)
        t = [torch.from_numpy(c[k]) for k in ("hm", "wh", "reg")]
        for stage, osz in (("raw", None), ("post", c["output_sizes"])):
            res = RL.ref_inference(t[0], t[1], t[2], c["image_sizes"], osz, **c["cfg"])
            counts = np.array([len(r) for r in res], np.int64)
            out["%s/%s/counts" % (name, stage)] = counts
            out["%s/%s/image_sizes" % (name, stage)] = np.array([list(r.image_size) for r in res], np.int64)
            out["%s/%s/boxes" % (name, stage)] = torch.cat([r.pred_boxes.tensor for r in res]).numpy()
            out["%s/%s/scores" % (name, stage)] = torch.cat([r.scores for r in res]).numpy()
            cls = torch.cat([r.pred_classes for r in res])
            assert cls.dtype == torch.int32
            out["%s/%s/classes" % (name, stage)] = cls.numpy()
        print(name, "raw", out[name + "/raw/counts"], "post", out[name + "/post/counts"])
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
