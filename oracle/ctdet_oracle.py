"""TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's ctdet hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (detectron2_centernet_b200) never does and fails
loudly when its CUDA library is missing.

Parity status: PINNED.  Every function here is checked (tests/test_oracle_golden.py) against
golden vectors produced by running the reference's own, unmodified functions in the build
container (oracle/ref_loader.py -> oracle/gen_golden.py -> tests/golden/*.npz), and directly
against the live reference whenever /root/reference is present.  The reference's own test-suite
has no golden vectors for this path.

Each function cites the reference file:line it restates (paths relative to the reference
root).  The restatement is written independently (vectorised numpy, explicit tie rule); it is
not a copy of the reference code.
"""
import math

import numpy as np
import torch

MAX_OBJS = 128          # hard-coded slot count, detectron2/data/detection_utils.py:611-616
BOX_DIVISOR = 4.0       # hard-coded "/ 4", detection_utils.py:618 (not down_ratio)


# --------------------------------------------------------------------------------------
# Target generation
# --------------------------------------------------------------------------------------
def gaussian_radius(height, width, min_overlap=0.7):
    """detection_utils.py:654-680.  Three quadratics, fp64, min of the larger roots
    (the third divides by 2 rather than 2a -- kept, it is what the reference computes)."""
    height = float(height)
    width = float(width)
    s = height + width
    area = width * height
    # case 1: a=1, b=s, c=area*(1-o)/(1+o)
    c1 = area * (1 - min_overlap) / (1 + min_overlap)
    r1 = (s + math.sqrt(s * s - 4.0 * c1)) / 2
    # case 2: a=4, b=2s, c=(1-o)*area
    b2 = 2 * s
    c2 = (1 - min_overlap) * width * height       # left-to-right, as the reference
    r2 = (b2 + math.sqrt(b2 * b2 - 16.0 * c2)) / 2
    # case 3: a=4o, b=-2o*s, c=(o-1)*area
    a3 = 4 * min_overlap
    b3 = -2 * min_overlap * s
    c3 = (min_overlap - 1) * width * height
    r3 = (b3 + math.sqrt(b3 * b3 - 4 * a3 * c3)) / 2
    return min(r1, r2, r3)


def _splat(plane, cx, cy, radius):
    """gaussian2D + draw_umich_gaussian, detection_utils.py:682-705: in-place max of a
    (2r+1)^2 fp64 Gaussian (sigma = (2r+1)/6), clipped to the map, rounded to fp32."""
    H, W = plane.shape
    if not (0 <= cx < W and 0 <= cy < H):
        # the reference's numpy slicing misbehaves (negative-index wrap / shape error) for
        # centres outside the map; boxes clipped to the image never produce them.  We skip.
        return
    x0, x1 = max(cx - radius, 0), min(cx + radius + 1, W)
    y0, y1 = max(cy - radius, 0), min(cy + radius + 1, H)
    sigma = (2 * radius + 1) / 6.0
    dx = np.arange(x0, x1, dtype=np.float64) - cx
    dy = np.arange(y0, y1, dtype=np.float64) - cy
    g = np.exp(-(dx[None, :] ** 2 + dy[:, None] ** 2) / (2 * sigma * sigma))
    g[g < np.finfo(np.float64).eps * 1.0] = 0     # :687, max of the kernel is exp(0) = 1
    win = plane[y0:y1, x0:x1]
    np.maximum(win, g.astype(np.float32), out=win)


def gen_heatmap(boxes, classes, H, W, C):
    """detection_utils.py:600-651 for one image.
    boxes f32 [n,4] XYXY in input pixels, classes int [n].  Returns dict of numpy arrays:
    hm f32[C,H,W], reg_mask u8[128], ind i64[128], wh f32[128,2], reg f32[128,2]."""
    boxes = np.asarray(boxes, dtype=np.float32).reshape(-1, 4)
    classes = np.asarray(classes, dtype=np.int64).reshape(-1)
    hm = np.zeros((C, H, W), np.float32)
    wh = np.zeros((MAX_OBJS, 2), np.float32)
    reg = np.zeros((MAX_OBJS, 2), np.float32)
    ind = np.zeros((MAX_OBJS,), np.int64)
    reg_mask = np.zeros((MAX_OBJS,), np.uint8)
    n = min(len(classes), MAX_OBJS)
    for k in range(n):
        b = boxes[k] / np.float32(BOX_DIVISOR)                 # :618 (fp32)
        bw = np.float32(b[2] - b[0])
        bh = np.float32(b[3] - b[1])
        if not (bh > 0 and bw > 0):                            # :621, slot k stays zero
            continue
        r = max(0, int(gaussian_radius(math.ceil(bh), math.ceil(bw))))   # :623-624
        ctx = np.float32(np.float32(b[0] + b[2]) / np.float32(2))       # :625-626 (fp32)
        cty = np.float32(np.float32(b[1] + b[3]) / np.float32(2))
        cxi, cyi = int(np.int32(ctx)), int(np.int32(cty))      # :627 trunc toward zero
        _splat(hm[int(classes[k])], cxi, cyi, r)               # :628
        wh[k] = (bw, bh)                                       # :629
        ind[k] = cyi * W + cxi                                 # :630
        reg[k] = (ctx - np.float32(cxi), cty - np.float32(cyi))  # :631
        reg_mask[k] = 1                                        # :632
    return {"hm": hm, "reg_mask": reg_mask, "ind": ind, "wh": wh, "reg": reg}


def gen_heatmap_batch(boxes_list, classes_list, H, W, C):
    """The torch.cat of per-image targets in CenterNet.losses, meta_arch/centernet.py:197-201."""
    per = [gen_heatmap(b, c, H, W, C) for b, c in zip(boxes_list, classes_list)]
    return {k: np.stack([p[k] for p in per]) for k in per[0]}


def count_positives(gt):
    """num_pos of _neg_loss (centernet.py:360): number of cells with gt == 1."""
    return int((gt["hm"] == 1).sum())


# --------------------------------------------------------------------------------------
# Losses (forward and backward), torch fp32 on CPU
# --------------------------------------------------------------------------------------
def _alpha_vector(alpha, C):
    """alpha handling of _neg_loss, centernet.py:341-349."""
    if alpha is None:
        return torch.ones(C)
    if not isinstance(alpha, (list, tuple)):
        return torch.full((C,), float(alpha))
    a = list(alpha)
    if len(a) == 1:
        a = a * C
    elif len(a) != C:
        a = a + [1] * (C - len(a))
    return torch.tensor(a, dtype=torch.float32)


def focal_loss(pred, gt, alpha=None):
    """_neg_loss, centernet.py:333-369.  pred = clamp(sigmoid(logits)) fp32 [B,C,H,W]."""
    a = _alpha_vector(alpha, pred.shape[1]).to(pred.dtype)
    is_pos = (gt == 1).to(pred.dtype)
    is_neg = (gt < 1).to(pred.dtype)
    pos_term = torch.log(pred) * (1 - pred) ** 2 * is_pos
    neg_term = torch.log(1 - pred) * pred ** 2 * (1 - gt) ** 4 * is_neg
    n_pos = is_pos.sum()
    pos_sum = (a[:, None, None] * pos_term.sum(0)).sum()      # alpha on positives only (:361)
    neg_sum = neg_term.sum()
    if n_pos == 0:
        return -neg_sum                                       # :365-366
    return -(pos_sum + neg_sum) / n_pos                       # :368


def gather_feat_nchw(feat, ind):
    """_transpose_and_gather_feat, centernet.py:383-397: feat[B,Ch,H,W], ind[B,K] ->
    [B,K,Ch] with out[b,k,c] = feat[b,c,ind[b,k]]  (no NHWC copy needed)."""
    B, Ch = feat.shape[:2]
    flat = feat.reshape(B, Ch, -1)
    idx = ind.long().unsqueeze(1).expand(B, Ch, ind.shape[1])
    return flat.gather(2, idx).permute(0, 2, 1)


def reg_l1_loss(output, mask, ind, target):
    """RegL1Loss.forward, centernet.py:376-381.  The mask is expanded over both channels
    before being summed, so the normaliser is 2*sum(mask) + 1e-4."""
    pred = gather_feat_nchw(output, ind)
    m = mask.unsqueeze(2).expand_as(pred).to(pred.dtype)
    return (pred * m - target * m).abs().sum() / (m.sum() + 1e-4)


def losses(hm_logits, wh, reg, gt, alpha=(1.0,), weights=(1.0, 0.1, 1.0)):
    """CenterNet.losses, centernet.py:191-212 (sigmoid + clamp(1e-4, 1-1e-4) then the three
    criteria and the HM/WH/OFF weights).  gt values may be numpy or torch."""
    g = {k: torch.as_tensor(v) for k, v in gt.items()}
    p = torch.clamp(torch.sigmoid(hm_logits), min=1e-4, max=1 - 1e-4)
    return {"hm_loss": focal_loss(p, g["hm"], list(alpha)) * weights[0],
            "wh_loss": reg_l1_loss(wh, g["reg_mask"], g["ind"], g["wh"]) * weights[1],
            "off_loss": reg_l1_loss(reg, g["reg_mask"], g["ind"], g["reg"]) * weights[2]}


def losses_and_grads(hm_logits, wh, reg, gt, alpha=(1.0,), weights=(1.0, 0.1, 1.0)):
    """losses() plus d(sum of the three weighted losses)/d(hm_logits, wh, reg) by autograd."""
    x = hm_logits.detach().clone().requires_grad_(True)
    w = wh.detach().clone().requires_grad_(True)
    r = reg.detach().clone().requires_grad_(True)
    out = losses(x, w, r, gt, alpha, weights)
    (out["hm_loss"] + out["wh_loss"] + out["off_loss"]).backward()
    return {k: v.detach() for k, v in out.items()}, (x.grad, w.grad, r.grad)


# --------------------------------------------------------------------------------------
# Decode
# --------------------------------------------------------------------------------------
def peak_mask(heat):
    """_nms, centernet.py:399-405: cell is kept iff it equals the max of its 3x3
    neighbourhood (-inf padding; plateaus are all kept)."""
    h = np.asarray(heat, dtype=np.float32)
    B, C, H, W = h.shape
    pad = np.full((B, C, H + 2, W + 2), -np.inf, np.float32)
    pad[:, :, 1:-1, 1:-1] = h
    m = h.copy()
    for dy in range(3):
        for dx in range(3):
            np.maximum(m, pad[:, :, dy:dy + H, dx:dx + W], out=m)
    return m == h


def decode(heat, wh, reg=None, down_ratio=1, K=100, cat_spec_wh=False, logits=False):
    """ctdet_decode, centernet.py:426-458, batched over images (the reference is batch=1).

    heat: f32 [B,C,H,W], probabilities in (0,1] as produced at centernet.py:161, or raw head
    logits when logits=True (sigmoid is monotone, so the ranking is taken on the logits and
    score = clamp(sigmoid(x), 1e-4, 1-1e-4) is applied to the K winners; the 3x3 peak test
    runs on the fp32 probabilities, exactly the values the reference compares).

    Selection: the reference takes a per-class top-K (:411) then a top-K of those C*K (:417);
    that equals one top-K over all C*H*W cells because no class can hold more than K of the
    global top K.  TIE RULE (torch.topk's is implementation-defined): higher score first, then
    lower flat index c*H*W + y*W + x.  Non-peak cells have score 0.0 (heat*keep, :405) and can
    only be selected when fewer than K peaks exist.

    Returns boxes f32[B,K,4], scores f32[B,K], classes i32[B,K], inds i64[B,K] (spatial)."""
    h = np.asarray(heat, dtype=np.float32)
    B, C, H, W = h.shape
    if H * W < K:
        raise RuntimeError("selected index k out of range")   # torch.topk at :411
    if logits:
        # the reference never sees logits: it ranks and peak-tests p = clamp(sigmoid(x)) in fp32 (:161, :401-404),
        # and distinct logits can share one fp32 probability (clamp saturation, rounding) -- then the reference
        # keeps every cell of such a plateau.  So the 3x3 test runs on p; the order among equal p is unspecified
        # in the reference (torch.topk), ours: higher logit first, then lower flat index.
        keep = peak_mask(_sigmoid_clamp(h))
        rank = np.where(keep, h, -np.inf)
    else:
        keep = peak_mask(h)
        rank = np.where(keep & (h > 0), h, np.float32(0))
    flat = rank.reshape(B, -1)
    order = np.stack([np.lexsort((np.arange(flat.shape[1]), -flat[b].astype(np.float64)))[:K]
                      for b in range(B)])
    top = np.take_along_axis(flat, order, 1)
    if logits:
        with np.errstate(over="ignore"):
            s = (np.float32(1) / (np.float32(1) + np.exp(-top.astype(np.float32)))).astype(np.float32)
        scores = np.where(np.isneginf(top), np.float32(0),
                          np.clip(s, np.float32(1e-4), np.float32(1 - 1e-4))).astype(np.float32)
    else:
        scores = top.astype(np.float32)
    classes = (order // (H * W)).astype(np.int32)              # :418
    inds = (order % (H * W)).astype(np.int64)
    ys = (inds // W).astype(np.float32)                        # :414
    xs = (inds % W).astype(np.float32)                         # :415
    whf = np.asarray(wh, np.float32).reshape(B, -1, H * W)
    if reg is not None:
        rf = np.asarray(reg, np.float32).reshape(B, 2, H * W)
        xs = xs + np.take_along_axis(rf[:, 0], inds, 1)        # :436-439
        ys = ys + np.take_along_axis(rf[:, 1], inds, 1)
    else:
        xs = xs + np.float32(0.5)                              # :440-441
        ys = ys + np.float32(0.5)
    if cat_spec_wh:                                            # :443-446 (wh has 2C channels)
        bidx = np.arange(B)[:, None]
        w_ = whf[bidx, 2 * classes, inds]
        h_ = whf[bidx, 2 * classes + 1, inds]
    else:
        w_ = np.take_along_axis(whf[:, 0], inds, 1)
        h_ = np.take_along_axis(whf[:, 1], inds, 1)
    two = np.float32(2)
    dr = np.float32(down_ratio)
    boxes = np.stack([(xs - w_ / two) * dr, (ys - h_ / two) * dr,
                      (xs + w_ / two) * dr, (ys + h_ / two) * dr], axis=2).astype(np.float32)
    return boxes, scores, classes, inds


def postprocess_single(boxes, scores, classes, max_dets=100, score_thresh=0.05):
    """CenterNet.inference_single_image, centernet.py:251-261: slice to max_dets then keep
    scores strictly greater than the threshold."""
    n = min(max_dets, len(scores))
    b, s, c = boxes[:n], scores[:n], classes[:n]
    k = s > np.float32(score_thresh)
    return b[k], s[k], c[k]


def detector_postprocess(boxes, scores, classes, image_size, output_size):
    """detector_postprocess, modeling/postprocessing.py:11-74, for the fields ctdet produces: the python-float
    quotients scale_x = out_w / in_w, scale_y = out_h / in_h (:39-42) multiply the fp32 box tensor in place
    (Boxes.scale, structures/boxes.py:271-276: the scalar is rounded to fp32 first), Boxes.clip (:184-197) clamps
    x to [0, out_w] and y to [0, out_h], and rows whose width or height is not > 0 are dropped (Boxes.nonempty
    :199-213, postprocessing.py:57).  image_size / output_size are (h, w)."""
    b = np.array(boxes, np.float32, copy=True).reshape(-1, 4)
    sx = np.float32(float(output_size[1]) / float(image_size[1]))
    sy = np.float32(float(output_size[0]) / float(image_size[0]))
    b[:, 0::2] *= sx
    b[:, 1::2] *= sy
    b[:, 0::2] = np.clip(b[:, 0::2], np.float32(0), np.float32(output_size[1]))
    b[:, 1::2] = np.clip(b[:, 1::2], np.float32(0), np.float32(output_size[0]))
    keep = ((b[:, 2] - b[:, 0]) > 0) & ((b[:, 3] - b[:, 1]) > 0)
    return b[keep], np.asarray(scores)[keep], np.asarray(classes)[keep]


def inference(hm_logits, wh, reg, image_sizes, output_sizes=None, down_ratio=4, topk_candidates=100,
              max_detections_per_image=100, score_threshold=0.05):
    """The eval branch from the head outputs on: sigmoid_ + clamp (centernet.py:161), `inference` (:214-234) with
    `inference_single_image` (:236-266) per image, then detector_postprocess (:169-171) when output_sizes is
    given.  -> list of (boxes f32[n,4], scores f32[n], classes i32[n])."""
    heat = _sigmoid_clamp(hm_logits)
    ob, os_, oc, _ = decode(heat, wh, reg, down_ratio, topk_candidates)
    res = []
    for i in range(heat.shape[0]):
        r = postprocess_single(ob[i], os_[i], oc[i], max_detections_per_image, score_threshold)
        if output_sizes is not None:
            r = detector_postprocess(r[0], r[1], r[2], image_sizes[i], output_sizes[i])
        res.append(r)
    return res


def _sigmoid_clamp(x):
    """clamp(sigmoid(x), 1e-4, 1 - 1e-4) in fp32 (centernet.py:161)."""
    x = np.asarray(x, np.float32)
    with np.errstate(over="ignore"):
        p = (np.float32(1) / (np.float32(1) + np.exp(-x))).astype(np.float32)
    return np.clip(p, np.float32(1e-4), np.float32(1 - 1e-4)).astype(np.float32)


def decode_torch(heat, wh, reg=None, down_ratio=1, K=100):
    """ctdet_decode restated with the reference's own operator sequence (max_pool2d, two
    torch.topk, gathers; centernet.py:399-458), batched.  Same results as decode() on tie-free
    inputs (checked in tests); this is the variant timed as the CPU baseline because it costs
    what the reference costs, whereas decode() pays for an explicit full sort."""
    heat = torch.as_tensor(heat)
    wh = torch.as_tensor(wh)
    B, C, H, W = heat.shape
    pooled = torch.nn.functional.max_pool2d(heat, 3, stride=1, padding=1)
    kept = heat * (pooled == heat).to(heat.dtype)
    cls_scores, cls_inds = torch.topk(kept.view(B, C, -1), K)            # per class (:411)
    scores, sel = torch.topk(cls_scores.view(B, -1), K)                  # over C*K (:417)
    classes = (sel // K).to(torch.int32)
    inds = cls_inds.view(B, -1).gather(1, sel)
    ys = (inds // W).to(torch.float32)
    xs = (inds % W).to(torch.float32)
    if reg is not None:
        r = gather_feat_nchw(torch.as_tensor(reg), inds)
        xs = xs + r[..., 0]
        ys = ys + r[..., 1]
    else:
        xs = xs + 0.5
        ys = ys + 0.5
    w = gather_feat_nchw(wh, inds)
    boxes = torch.stack([xs - w[..., 0] / 2, ys - w[..., 1] / 2,
                         xs + w[..., 0] / 2, ys + w[..., 1] / 2], dim=2) * down_ratio
    return boxes.numpy(), scores.numpy(), classes.numpy(), inds.numpy()


# --------------------------------------------------------------------------------------
# Input pre-processing
def preprocess_image(images, pixel_mean, pixel_std, size_divisibility=0):
    """CenterNet.preprocess_image, meta_arch/centernet.py:177-184: x / 255., (x - mean) / std, then
    ImageList.from_tensors (structures/image_list.py:58-130): zero padding (pad_value 0.0) at the
    bottom / right up to the batch maximum rounded up to size_divisibility.
    images: list of torch CPU tensors [C,h,w] (uint8 or float32).  Returns (tensor f32 [N,C,Hp,Wp],
    [(h, w), ...])."""
    mean = torch.tensor(list(pixel_mean), dtype=torch.float32).view(-1, 1, 1)       # :64
    std = torch.tensor(list(pixel_std), dtype=torch.float32).view(-1, 1, 1)         # :65
    xs = [x / 255. for x in images]                                                   # :177
    xs = [(x - mean) / std for x in xs]                                               # :183
    sizes = [tuple(int(v) for v in x.shape[-2:]) for x in xs]                         # image_list.py:111
    Hp, Wp = max(s[0] for s in sizes), max(s[1] for s in sizes)                       # :83-97
    if size_divisibility > 1:                                                         # :99-102
        Hp = (Hp + size_divisibility - 1) // size_divisibility * size_divisibility
        Wp = (Wp + size_divisibility - 1) // size_divisibility * size_divisibility
    out = xs[0].new_full((len(xs), xs[0].shape[0], Hp, Wp), 0.0)                      # :124-125
    for x, dst in zip(xs, out):
        dst[..., :x.shape[-2], :x.shape[-1]].copy_(x)                                 # :126-127
    return out, sizes
