"""TEST / MEASUREMENT INFRASTRUCTURE ONLY -- the reference's own OPERATOR SEQUENCE for the ctdet hot path, device
agnostic, so that bench.py can time "the reference on the B200" (SURVEY.md section 8d "Timing method", BASELINE.md
section 3): the same ATen kernels in the same order the reference launches them, on CUDA tensors.

oracle/ctdet_oracle.py restates WHAT the reference computes (vectorised, explicit tie rule) and is what parity is
checked against; this file restates HOW it computes it -- every full-map pass, permute copy and per-image Python loop
that costs the reference its time -- and is pinned on the oracle by tests/test_ref_gpu_chain.py (CPU).  The
`detectron2` package cannot be imported on the GPU box (no /root/reference there, and fvcore / yacs / pycocotools /
detectron2._C are absent everywhere), so the reference's functions themselves cannot be timed there.

  target_gen_host   detectron2/data/detection_utils.py:600-651 per image on the HOST (numpy), then the `.to(device)` +
                    torch.cat of meta_arch/centernet.py:197-201  (kind: port = oracle.gen_heatmap)
  losses            centernet.py:191-212: sigmoid_ + clamp, _neg_loss :333-369 op by op, RegL1Loss :372-381 with the
                    full NHWC permute of _transpose_and_gather_feat :383-387, autograd backward
  decode_loop       centernet.py:214-234 per-image Python loop of ctdet_decode :426-458 (_nms :399-405, _topk :408-424)
"""
import torch
import torch.nn.functional as F

from . import ctdet_oracle as O


def _permute_gather(feat, ind):
    """centernet.py:383-397: NHWC permute + contiguous copy of the whole map, then gather of K rows."""
    feat = feat.permute(0, 2, 3, 1).contiguous()
    feat = feat.view(feat.size(0), -1, feat.size(3))
    idx = ind.unsqueeze(2).expand(ind.size(0), ind.size(1), feat.size(2))
    return feat.gather(1, idx)


def neg_loss(pred, gt, alpha):
    """centernet.py:333-369, one ATen op per reference op (each a full pass over the map)."""
    C = pred.shape[1]
    a = list(alpha) if isinstance(alpha, (list, tuple)) else [alpha]
    if len(a) == 1:
        a = a * C
    a = torch.tensor(a, dtype=torch.float32).to(pred.device)          # `.cuda()` at :342-349: an H2D per call
    pos_inds = gt.eq(1).float()
    neg_inds = gt.lt(1).float()
    neg_weights = torch.pow(1 - gt, 4)
    pos_loss = torch.log(pred) * torch.pow(1 - pred, 2) * pos_inds
    neg = torch.log(1 - pred) * torch.pow(pred, 2) * neg_weights * neg_inds
    num_pos = pos_inds.float().sum()
    pos_loss = (a[:, None, None] * pos_loss.sum(0)).sum()
    neg = neg.sum()
    if num_pos == 0:                                                   # host sync (:365)
        return 0 - neg
    return 0 - (pos_loss + neg) / num_pos


def reg_l1(output, mask, ind, target):
    """centernet.py:376-381."""
    pred = _permute_gather(output, ind)
    m = mask.unsqueeze(2).expand_as(pred).float()
    return F.l1_loss(pred * m, target * m, reduction="sum") / (m.sum() + 1e-4)


def losses_fwd_bwd(hm_logits, wh, reg, gt, alpha=(1.0,), weights=(1.0, 0.1, 1.0)):
    """centernet.py:191-212 + backward of the summed losses.  hm_logits / wh / reg: leaf tensors with
    requires_grad; gt: dict of device tensors.  Returns the loss dict (grads land in .grad)."""
    hm = (hm_logits * 1).sigmoid_()                                    # the head output is a non-leaf in the reference
    out = {"hm_loss": neg_loss(torch.clamp(hm, min=1e-4, max=1 - 1e-4), gt["hm"], alpha) * weights[0],
           "wh_loss": reg_l1(wh, gt["reg_mask"], gt["ind"], gt["wh"]) * weights[1],
           "off_loss": reg_l1(reg, gt["reg_mask"], gt["ind"], gt["reg"]) * weights[2]}
    (out["hm_loss"] + out["wh_loss"] + out["off_loss"]).backward()
    return out


def target_gen_host(boxes_list, classes_list, H, W, C, device):
    """detection_utils.py:600-651 per image on the host, then centernet.py:197-201."""
    per = [O.gen_heatmap(b, c, H, W, C) for b, c in zip(boxes_list, classes_list)]
    per = [{k: torch.tensor(v) for k, v in p.items()} for p in per]    # `torch.tensor(hm)` copies (:644-649)
    return {k: torch.cat([p[k].unsqueeze(0).to(device) for p in per], dim=0) for k in per[0]}


def _nms(heat):
    hmax = F.max_pool2d(heat, (3, 3), stride=1, padding=1)
    keep = (hmax == heat).float()
    return heat * keep


def _topk(scores, K):
    batch, cat, height, width = scores.size()
    topk_scores, topk_inds = torch.topk(scores.view(batch, cat, -1), K)
    topk_inds = topk_inds % (height * width)
    topk_ys = (topk_inds // width).int().float()
    topk_xs = (topk_inds % width).int().float()
    topk_score, topk_ind = torch.topk(topk_scores.view(batch, -1), K)
    topk_clses = (topk_ind // K).int()

    def g(t):
        return t.view(batch, -1, 1).gather(1, topk_ind.unsqueeze(2)).view(batch, K)
    return topk_score, g(topk_inds), topk_clses, g(topk_ys), g(topk_xs)


def ctdet_decode_one(heat, wh, reg, down_ratio, K):
    """centernet.py:426-458 for a batch of one."""
    heat = _nms(heat)
    scores, inds, clses, ys, xs = _topk(heat, K)
    r = _permute_gather(reg, inds).view(1, K, 2)
    xs = xs.view(1, K, 1) + r[:, :, 0:1]
    ys = ys.view(1, K, 1) + r[:, :, 1:2]
    w = _permute_gather(wh, inds).view(1, K, 2)
    boxes = torch.cat([xs - w[..., 0:1] / 2, ys - w[..., 1:2] / 2,
                       xs + w[..., 0:1] / 2, ys + w[..., 1:2] / 2], dim=2) * down_ratio
    return boxes.view(K, 4), scores.view(K), clses.view(K)


def decode_loop(hm_logits, wh, reg, down_ratio=4, K=100):
    """The eval branch: sigmoid_ + clamp on the whole batch (:161), then the per-image loop of `inference` (:224-233)."""
    heat = torch.clamp(hm_logits.clone().sigmoid_(), min=1e-4, max=1 - 1e-4)
    outs = [ctdet_decode_one(heat[i].unsqueeze(0), wh[i].unsqueeze(0), reg[i].unsqueeze(0), down_ratio, K)
            for i in range(heat.shape[0])]
    return (torch.stack([o[0] for o in outs]), torch.stack([o[1] for o in outs]), torch.stack([o[2] for o in outs]))
