"""ctdet losses on the GPU -- the operator surface of detectron2/modeling/meta_arch/centernet.py:
`FocalLoss` (:323), `_neg_loss` (:333), `RegL1Loss` (:372) and the fused `CenterNet.losses`
(:191-212), backed by the single-pass forward+backward kernel `ctdet_loss_fwd_bwd`.
"""
import ctypes

import torch
import torch.nn as nn

from . import _lib

_FLAG_PROB = 1
_FLAG_ALPHA_UNIFORM = 2


_alpha_cache = {}


def alpha_is_uniform(alpha):
    """True when every class shares one alpha (None, a scalar, or a 1-element list)."""
    if alpha is None or isinstance(alpha, (int, float)):
        return True
    if isinstance(alpha, torch.Tensor):
        return alpha.numel() == 1
    return len(alpha) == 1


def alpha_tensor(alpha, num_classes, device):
    """alpha handling of _neg_loss (centernet.py:341-349) -> f32[C] on `device`.  Unlike the
    reference the caller's list is not mutated (`alpha.extend` at :348) and the tensor is built and
    copied to the device once per distinct value (the reference does an H2D copy every call)."""
    if not isinstance(alpha, torch.Tensor):
        key = (None if alpha is None else (tuple(alpha) if isinstance(alpha, (list, tuple)) else float(alpha)),
               num_classes, str(device))
        hit = _alpha_cache.get(key)
        if hit is None:
            hit = _alpha_cache[key] = _alpha_tensor(alpha, num_classes, device)
        return hit
    return _alpha_tensor(alpha, num_classes, device)


def _alpha_tensor(alpha, num_classes, device):
    if isinstance(alpha, torch.Tensor):
        a = alpha.to(device=device, dtype=torch.float32).reshape(-1)
        if a.numel() == 1:
            a = a.expand(num_classes)
        return a.contiguous()
    if alpha is None:
        vals = [1.0] * num_classes
    elif not isinstance(alpha, (list, tuple)):
        vals = [float(alpha)] * num_classes
    else:
        vals = [float(v) for v in alpha]
        if len(vals) == 1:
            vals = vals * num_classes
        elif len(vals) != num_classes:
            vals = vals + [1.0] * (num_classes - len(vals))
    return torch.tensor(vals, dtype=torch.float32).to(device, non_blocking=True)


def _call(hm, wh, reg, gt_hm, reg_mask, ind, gt_wh, gt_reg, alpha, num_pos, norm, flags, weights,
          need_grad):
    lib = _lib.load()
    ref = hm if hm is not None else (wh if wh is not None else reg)
    dev = ref.device
    if hm is not None:
        B, C, H, W = hm.shape
    else:
        B, _, H, W = ref.shape
        C = 1
    M = ind.shape[1] if ind is not None else 1
    losses = torch.empty(3, dtype=torch.float32, device=dev)
    d_hm = torch.empty_like(hm) if (need_grad and hm is not None) else None
    d_wh = torch.empty_like(wh) if (need_grad and wh is not None) else None
    d_reg = torch.empty_like(reg) if (need_grad and reg is not None) else None
    nbytes = lib.ctdet_loss_workspace_bytes()
    ws = _lib.zero_workspace("loss", nbytes, dev)
    with torch.cuda.device(dev):
        rc = lib.ctdet_loss_fwd_bwd(
            _lib.ptr(hm), _lib.ptr(wh), _lib.ptr(reg), _lib.ptr(gt_hm), _lib.ptr(reg_mask),
            _lib.ptr(ind), _lib.ptr(gt_wh), _lib.ptr(gt_reg), _lib.ptr(alpha), _lib.ptr(num_pos),
            _lib.ptr(norm), B, C, H, W, M, flags, weights[0], weights[1], weights[2],
            _lib.ptr(losses), _lib.ptr(d_hm), _lib.ptr(d_wh), _lib.ptr(d_reg),
            _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(dev))
    _lib.check(rc, "ctdet_loss_fwd_bwd")
    return losses, d_hm, d_wh, d_reg


def _scaled(d, g):
    """d *= g on the device; the kernel exits immediately when g == 1 (no host sync)."""
    if d is None:
        return None
    lib = _lib.load()
    g = g.to(dtype=torch.float32).contiguous()
    with torch.cuda.device(d.device):
        rc = lib.ctdet_scale_grad(_lib.ptr(d), ctypes.c_size_t(d.numel()), _lib.ptr(g),
                                  _lib.stream_ptr(d.device))
    _lib.check(rc, "ctdet_scale_grad")
    return d


class _FusedLoss(torch.autograd.Function):
    """losses[3] = (hm, wh, off) * weights; gradients are produced by the same forward sweep and
    only rescaled in backward if the upstream gradient differs from 1."""

    @staticmethod
    def forward(ctx, hm, wh, reg, gt_hm, reg_mask, ind, gt_wh, gt_reg, alpha, num_pos, norm, flags,
                weights):
        need = any(t is not None and t.requires_grad for t in (hm, wh, reg))
        losses, d_hm, d_wh, d_reg = _call(hm, wh, reg, gt_hm, reg_mask, ind, gt_wh, gt_reg, alpha,
                                          num_pos, norm, flags, weights, need)
        ctx.grads = (d_hm, d_wh, d_reg)
        return losses

    @staticmethod
    def backward(ctx, g):
        if ctx.grads is None:
            raise RuntimeError("the fused ctdet losses were already back-propagated: the gradients of the three losses are "
                               "produced by ONE forward sweep and released by the first backward -- sum hm_loss, wh_loss "
                               "and off_loss before calling backward once (retain_graph / per-loss backward is not supported)")
        d_hm, d_wh, d_reg = ctx.grads
        ctx.grads = None
        g = g.contiguous()
        return (_scaled(d_hm, g[0:1]), _scaled(d_wh, g[1:2]), _scaled(d_reg, g[2:3]),
                None, None, None, None, None, None, None, None, None, None)


def _prep(t, what):
    if t is None:
        return None
    _lib.require_cuda(what, t)
    return _lib.f32c(t, what)


def ctdet_losses(outputs, targets, alpha=None, weights=(1.0, 0.1, 1.0), norm=None):
    """Fused CenterNet.losses (centernet.py:191-212).

    outputs: {'hm': logits [B,C,H,W], 'wh': [B,2,H,W], 'reg': [B,2,H,W]} -- the raw head outputs
    (sigmoid + clamp of :204 happen inside; unlike the reference `outputs['hm']` is NOT modified in
    place).  targets: dict hm/reg_mask/ind/wh/reg (+ optional num_pos i32[B]) as produced by
    gen_heatmap_batch.  norm: optional f32[2] device tensor {num_pos, sum(reg_mask)} overriding the
    local normalisers (used for the cross-rank all-reduce).  Returns the reference's dict."""
    hm = _prep(outputs["hm"], "ctdet losses")
    wh = _prep(outputs["wh"], "ctdet losses")
    reg = _prep(outputs["reg"], "ctdet losses")
    gt_hm = _prep(targets["hm"], "ctdet losses")
    if gt_hm.shape != hm.shape:
        raise RuntimeError("gt hm shape %s != hm shape %s" % (tuple(gt_hm.shape), tuple(hm.shape)))
    a = alpha_tensor(alpha, hm.shape[1], hm.device)
    out = _FusedLoss.apply(hm, wh, reg, gt_hm, targets["reg_mask"].contiguous(),
                           targets["ind"].contiguous(), _prep(targets["wh"], "gt wh"),
                           _prep(targets["reg"], "gt reg"), a, targets.get("num_pos"), norm,
                           _FLAG_ALPHA_UNIFORM if alpha_is_uniform(alpha) else 0,
                           tuple(float(w) for w in weights))
    return {"hm_loss": out[0], "wh_loss": out[1], "off_loss": out[2]}


class _FusedTargetLoss(torch.autograd.Function):
    """ctdet_target_loss_fused: targets from the objects and the three losses in one pass; the
    gradients come out of the same sweep (see _FusedLoss)."""

    @staticmethod
    def forward(ctx, hm, wh, reg, boxes, classes, offsets, alpha, norm, flags, weights, slots):
        lib = _lib.load()
        dev = hm.device
        B, C, H, W = hm.shape
        need = any(t.requires_grad for t in (hm, wh, reg))
        losses = torch.empty(3, dtype=torch.float32, device=dev)
        d_hm = torch.empty_like(hm) if need else None
        d_wh = torch.empty_like(wh) if need else None
        d_reg = torch.empty_like(reg) if need else None
        ws = _lib.zero_workspace("loss_fused", lib.ctdet_fused_workspace_bytes(B, C), dev)
        with torch.cuda.device(dev):
            rc = lib.ctdet_target_loss_fused(
                _lib.ptr(boxes), _lib.ptr(classes), _lib.ptr(offsets), _lib.ptr(hm), _lib.ptr(wh), _lib.ptr(reg),
                _lib.ptr(alpha), _lib.ptr(norm), B, C, H, W, flags, weights[0], weights[1], weights[2],
                _lib.ptr(slots["reg_mask"]), _lib.ptr(slots["ind"]), _lib.ptr(slots["wh"]), _lib.ptr(slots["reg"]),
                _lib.ptr(slots["num_pos"]), _lib.ptr(losses), _lib.ptr(d_hm), _lib.ptr(d_wh), _lib.ptr(d_reg),
                _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(dev))
        _lib.check(rc, "ctdet_target_loss_fused")
        ctx.grads = (d_hm, d_wh, d_reg)
        return losses

    @staticmethod
    def backward(ctx, g):
        if ctx.grads is None:
            raise RuntimeError("the fused ctdet losses were already back-propagated: the gradients of the three losses are "
                               "produced by ONE forward sweep and released by the first backward -- sum hm_loss, wh_loss "
                               "and off_loss before calling backward once (retain_graph / per-loss backward is not supported)")
        d_hm, d_wh, d_reg = ctx.grads
        ctx.grads = None
        g = g.contiguous()
        return (_scaled(d_hm, g[0:1]), _scaled(d_wh, g[1:2]), _scaled(d_reg, g[2:3]),
                None, None, None, None, None, None, None, None)


def ctdet_losses_from_objects(outputs, boxes, classes, offsets, alpha=None, weights=(1.0, 0.1, 1.0), norm=None,
                              slots=None):
    """gen_heatmap for the batch (detection_utils.py:600, called at centernet.py:188) fused with
    CenterNet.losses (:191-212): the heatmap target is generated band by band in shared memory and
    never written to HBM.  Objects in the CSR form of `target.pack_instances` / ctdet_target_gen:
    boxes f32 [total,4] XYXY input pixels, classes i64 [total], offsets i32 [B+1] (device tensors).
    Returns (loss dict, slot targets dict {reg_mask, ind, wh, reg, num_pos}); `slots` lets the
    caller reuse the slot buffers between steps."""
    hm = _prep(outputs["hm"], "ctdet losses")
    wh = _prep(outputs["wh"], "ctdet losses")
    reg = _prep(outputs["reg"], "ctdet losses")
    _lib.require_cuda("ctdet losses", boxes, classes, offsets)
    B, C = hm.shape[0], hm.shape[1]
    dev = hm.device
    if offsets.numel() != B + 1:
        raise RuntimeError("offsets must have B + 1 = %d entries, got %d" % (B + 1, offsets.numel()))
    if slots is None:
        slots = {"reg_mask": torch.empty((B, 128), dtype=torch.uint8, device=dev),
                 "ind": torch.empty((B, 128), dtype=torch.int64, device=dev),
                 "wh": torch.empty((B, 128, 2), dtype=torch.float32, device=dev),
                 "reg": torch.empty((B, 128, 2), dtype=torch.float32, device=dev),
                 "num_pos": torch.empty((B,), dtype=torch.int32, device=dev)}
    a = alpha_tensor(alpha, C, dev)
    out = _FusedTargetLoss.apply(hm, wh, reg, boxes.contiguous(), classes.contiguous(), offsets.contiguous(), a, norm,
                                 _FLAG_ALPHA_UNIFORM if alpha_is_uniform(alpha) else 0,
                                 tuple(float(w) for w in weights), slots)
    return {"hm_loss": out[0], "wh_loss": out[1], "off_loss": out[2]}, slots


def _neg_loss(pred, gt, alpha=None, num_pos=None):
    """Drop-in for centernet.py:333: pred = clamp(sigmoid(.)) probabilities [B,C,H,W]."""
    pred = _prep(pred, "_neg_loss")
    gt = _prep(gt, "_neg_loss")
    a = alpha_tensor(alpha, pred.shape[1], pred.device)
    out = _FusedLoss.apply(pred, None, None, gt, None, None, None, None, a, num_pos, None,
                           _FLAG_PROB | (_FLAG_ALPHA_UNIFORM if alpha_is_uniform(alpha) else 0),
                           (1.0, 0.0, 0.0))
    return out[0]


class FocalLoss(nn.Module):
    """nn.Module wrapper for the focal loss, centernet.py:323-331."""

    def __init__(self, alpha):
        super().__init__()
        self.neg_loss = _neg_loss
        self.alpha = alpha

    def forward(self, out, target):
        return self.neg_loss(out, target, self.alpha)


class RegL1Loss(nn.Module):
    """centernet.py:372-381: output [B,2,H,W], mask u8/bool [B,M], ind i64 [B,M], target [B,M,2]."""

    def forward(self, output, mask, ind, target):
        output = _prep(output, "RegL1Loss")
        if output.shape[1] != 2:
            raise NotImplementedError("RegL1Loss kernel handles the 2-channel wh/reg heads")
        mask = mask.to(torch.uint8).contiguous()
        out = _FusedLoss.apply(None, output, None, None, mask, ind.contiguous(),
                               _prep(target, "RegL1Loss target"), None, None, None, None, 0,
                               (0.0, 1.0, 0.0))
        return out[1]


# ---------------------------------------------------------------------------------------------------------------
# SURVEY section 8 f2, second half: hm head 1x1 convolution + focal loss in one kernel
def _head_focal_call(feat, weight, bias, gt_hm, alpha, num_pos, norm, uniform, w_hm, need_grad):
    lib = _lib.load()
    dev = feat.device
    B, Cf, H, W = feat.shape
    C = weight.shape[0]
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    d_logits = torch.empty((B, C, H, W), dtype=torch.float32, device=dev) if need_grad else None
    ws = _lib.zero_workspace("loss", lib.ctdet_loss_workspace_bytes(), dev)
    with torch.cuda.device(dev):
        rc = lib.ctdet_head_focal_fused(_lib.ptr(feat), _lib.ptr(weight), _lib.ptr(bias), _lib.ptr(gt_hm), _lib.ptr(alpha),
                                        _lib.ptr(num_pos), _lib.ptr(norm), B, Cf, C, H, W, _FLAG_ALPHA_UNIFORM if uniform else 0,
                                        float(w_hm), _lib.ptr(loss), _lib.ptr(d_logits), _lib.ptr(ws),
                                        ctypes.c_size_t(ws.numel()), _lib.stream_ptr(dev))
    _lib.check(rc, "ctdet_head_focal_fused")
    return loss, d_logits


def _head_focal_bwd_call(feat, weight, bias, gt_hm, alpha, num_pos, norm, uniform, w_hm, want_d_logits=False):
    """The kernel with the convolution's backward inside: -> loss, d_feat, d_weight, d_bias (, d_logits)."""
    lib = _lib.load()
    dev = feat.device
    B, Cf, H, W = feat.shape
    C = weight.shape[0]
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    d_feat = torch.empty_like(feat)
    d_weight = torch.empty((C, Cf), dtype=torch.float32, device=dev)
    d_bias = torch.empty((C,), dtype=torch.float32, device=dev)
    d_logits = torch.empty((B, C, H, W), dtype=torch.float32, device=dev) if want_d_logits else None
    ws = _lib.zero_workspace("loss", lib.ctdet_loss_workspace_bytes(), dev)
    with torch.cuda.device(dev):
        rc = lib.ctdet_head_focal_fused_bwd(_lib.ptr(feat), _lib.ptr(weight), _lib.ptr(bias), _lib.ptr(gt_hm), _lib.ptr(alpha),
                                            _lib.ptr(num_pos), _lib.ptr(norm), B, Cf, C, H, W,
                                            _FLAG_ALPHA_UNIFORM if uniform else 0, float(w_hm), _lib.ptr(loss), _lib.ptr(d_logits),
                                            _lib.ptr(d_feat), _lib.ptr(d_weight), _lib.ptr(d_bias), _lib.ptr(ws),
                                            ctypes.c_size_t(ws.numel()), _lib.stream_ptr(dev))
    _lib.check(rc, "ctdet_head_focal_fused_bwd")
    return loss, d_feat, d_weight, d_bias, d_logits


def head_bwd_fusable(Cf, C, HW):
    """Shapes `ctdet_head_focal_fused_bwd` tiles (TMEM: logits + dW^T + d_feat <= 512 columns)."""
    return Cf % 128 == 0 and Cf <= 256 and C % 16 == 0 and C + (Cf // 128) * C + Cf <= 512


class _HeadFocalFused(torch.autograd.Function):
    """hm_loss * w_hm of `Conv2d(Cf, C, 1)(feat)` without the logits in HBM.  Where the shapes allow it the convolution's
    backward (grad_input / grad_weight / grad_bias) is produced by the same kernel -- the gradient of the logits never
    reaches HBM either; otherwise the kernel returns d(loss)/d(logits) and two torch GEMMs finish the job."""

    @staticmethod
    def forward(ctx, feat, weight, bias, gt_hm, alpha, num_pos, norm, uniform, w_hm, fuse_backward):
        need = feat.requires_grad or weight.requires_grad or (bias is not None and bias.requires_grad)
        ctx.has_bias = bias is not None
        ctx.fused = bool(need and fuse_backward and head_bwd_fusable(feat.shape[1], weight.shape[0], feat.shape[2] * feat.shape[3]))
        if ctx.fused:
            loss, d_feat, d_weight, d_bias, _ = _head_focal_bwd_call(feat, weight, bias, gt_hm, alpha, num_pos, norm, uniform, w_hm)
            ctx.grads = (d_feat, d_weight, d_bias)
            return loss
        loss, d_logits = _head_focal_call(feat, weight, bias, gt_hm, alpha, num_pos, norm, uniform, w_hm, need)
        ctx.save_for_backward(feat, weight)
        ctx.grads = d_logits
        return loss

    @staticmethod
    def backward(ctx, g):
        if ctx.grads is None:
            raise RuntimeError("the fused hm head loss was already back-propagated (or no input required a gradient)")
        g = g.contiguous()[0:1]
        if ctx.fused:
            d_feat, d_weight, d_bias = ctx.grads
            ctx.grads = None
            return (_scaled(d_feat, g) if ctx.needs_input_grad[0] else None,
                    _scaled(d_weight, g) if ctx.needs_input_grad[1] else None,
                    _scaled(d_bias, g) if (ctx.has_bias and ctx.needs_input_grad[2]) else None,
                    None, None, None, None, None, None, None)
        feat, weight = ctx.saved_tensors
        d = _scaled(ctx.grads, g)
        ctx.grads = None
        B, C = d.shape[0], d.shape[1]
        d2 = d.view(B, C, -1)                                     # [B, C, P]
        f2 = feat.view(B, feat.shape[1], -1)                      # [B, Cf, P]
        g_feat = torch.matmul(weight.t().unsqueeze(0), d2).view_as(feat) if ctx.needs_input_grad[0] else None
        g_w = torch.einsum("bcp,bfp->cf", d2, f2) if ctx.needs_input_grad[1] else None
        g_b = d2.sum(dim=(0, 2)) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        return g_feat, g_w, g_b, None, None, None, None, None, None, None


def hm_head_focal_loss(feat, weight, bias, gt_hm, alpha=None, num_pos=None, norm=None, w_hm=1.0, fuse_backward=True):
    """`FocalLoss()(clamp(sigmoid_(Conv2d(Cf, C, 1)(feat))), gt_hm) * w_hm` (centernet.py:118-123, :152, :204, :323-369)
    in one kernel: the logits are formed per pixel tile on the tensor cores (bf16 operands) and never written to HBM.
    feat f32 [B,Cf,H,W]; weight f32 [C,Cf] or [C,Cf,1,1]; bias f32 [C] or None; gt_hm f32 [B,C,H,W];
    num_pos i32 [B] (from generate_targets) or norm f32 [2] = {num_pos, .}; both None: positives are counted."""
    feat = _prep(feat, "feat")
    gt_hm = _prep(gt_hm, "gt_hm")
    if weight.dim() == 4:
        if weight.shape[2:] != (1, 1):
            raise NotImplementedError("hm_head_focal_loss fuses the head's final 1x1 convolution")
        weight = weight.reshape(weight.shape[0], weight.shape[1])
    weight = _prep(weight, "weight")
    bias = _prep(bias, "bias")
    C = weight.shape[0]
    if tuple(gt_hm.shape) != (feat.shape[0], C, feat.shape[2], feat.shape[3]):
        raise RuntimeError("gt_hm must be [%d,%d,%d,%d]" % (feat.shape[0], C, feat.shape[2], feat.shape[3]))
    a = alpha_tensor(alpha, C, feat.device)
    return _HeadFocalFused.apply(feat, weight, bias, gt_hm, a, num_pos, norm, alpha_is_uniform(alpha), w_hm, fuse_backward)


def ctdet_losses_fused_head(feat, weight, bias, outputs, targets, alpha=None, weights=(1.0, 0.1, 1.0), norm=None,
                            fuse_backward=True):
    """CenterNet.losses (centernet.py:191-212) with the 'hm' branch's final 1x1 convolution folded in: `feat` is the
    input of that convolution (the ReLU output of centernet.py:118-123), `weight` / `bias` its parameters, and
    `outputs` holds only 'wh' and 'reg'.  hm_loss comes from hm_head_focal_loss (the logits never reach HBM), wh_loss /
    off_loss from the RegL1 sections of ctdet_loss_fwd_bwd (hm == NULL).  Returns the reference's dict; as with
    ctdet_losses, sum the three before calling backward once."""
    if "hm" in outputs and outputs["hm"] is not None:
        raise RuntimeError("ctdet_losses_fused_head forms the hm logits itself: pass the head's input as `feat`, not "
                           "outputs['hm'] (use ctdet_losses for materialised logits)")
    wh = _prep(outputs["wh"], "ctdet losses")
    reg = _prep(outputs["reg"], "ctdet losses")
    num_pos = targets.get("num_pos")
    hm_loss = hm_head_focal_loss(feat, weight, bias, targets["hm"], alpha=alpha, num_pos=num_pos, norm=norm,
                                 w_hm=float(weights[0]), fuse_backward=fuse_backward)
    a = alpha_tensor(alpha, 1, wh.device)
    out = _FusedLoss.apply(None, wh, reg, None, targets["reg_mask"].contiguous(), targets["ind"].contiguous(),
                           _prep(targets["wh"], "gt wh"), _prep(targets["reg"], "gt reg"), a, num_pos, norm, 0,
                           (0.0, float(weights[1]), float(weights[2])))
    return {"hm_loss": hm_loss.reshape(()), "wh_loss": out[1], "off_loss": out[2]}
