"""ctdet decode on the GPU -- the operator surface of detectron2/modeling/meta_arch/centernet.py:
`_nms` (:399), `_topk` (:408), `_gather_feat` (:389), `_transpose_and_gather_feat` (:383) and
`ctdet_decode` (:426), backed by `ctdet_decode` / `ctdet_nms` / `ctdet_gather_feat`.

Tie rule (torch.topk's is implementation-defined): higher score first, then lower flat index
c*H*W + y*W + x.  On tie-free inputs the results equal the reference's bit for bit.
"""
import ctypes

import torch

from . import _lib

FLAG_LOGITS, FLAG_CAT_SPEC_WH, FLAG_NO_NMS = 1, 2, 4

def _workspace(nbytes, device, B):
    # zero-initialised once; the finalize kernel leaves the per-image headers zeroed.  The header
    # size depends on B, so each batch size gets its own buffer.
    return _lib.zero_workspace("decode%d" % B, nbytes, device)


def ctdet_decode_batch(heat, wh, reg=None, down_ratio=1, cat_spec_wh=False, K=100, logits=False,
                       nms=True):
    """Batched decode.  heat [B,C,H,W] probabilities (or raw logits with logits=True), wh
    [B,2,H,W] (or [B,2C,H,W] with cat_spec_wh), reg [B,2,H,W] or None.
    Returns boxes f32[B,K,4], scores f32[B,K] (descending), classes i32[B,K], inds i64[B,K]."""
    _lib.require_cuda("ctdet_decode", heat, wh, reg)
    lib = _lib.load()
    heat = _lib.f32c(heat, "heat")
    wh = _lib.f32c(wh, "wh")
    reg = None if reg is None else _lib.f32c(reg, "reg")
    if heat.dim() != 4:
        raise RuntimeError("heat must be [B,C,H,W]")
    B, C, H, W = heat.shape
    want_wh = 2 * C if cat_spec_wh else 2
    if tuple(wh.shape) != (B, want_wh, H, W):
        raise RuntimeError("wh must be [%d,%d,%d,%d], got %s" % (B, want_wh, H, W, tuple(wh.shape)))
    if reg is not None and tuple(reg.shape) != (B, 2, H, W):
        raise RuntimeError("reg must be [%d,2,%d,%d]" % (B, H, W))
    if K > H * W:
        raise RuntimeError("selected index k out of range")      # what torch.topk raises at :411
    dev = heat.device
    boxes = torch.empty((B, K, 4), dtype=torch.float32, device=dev)
    scores = torch.empty((B, K), dtype=torch.float32, device=dev)
    classes = torch.empty((B, K), dtype=torch.int32, device=dev)
    inds = torch.empty((B, K), dtype=torch.int64, device=dev)
    flags = (FLAG_LOGITS if logits else 0) | (FLAG_CAT_SPEC_WH if cat_spec_wh else 0) | \
            (0 if nms else FLAG_NO_NMS)
    nbytes = lib.ctdet_decode_workspace_bytes(B, C, H, W, K)
    ws = _workspace(max(nbytes, 8), dev, B)
    with torch.cuda.device(dev):
        rc = lib.ctdet_decode(_lib.ptr(heat), _lib.ptr(wh), _lib.ptr(reg), B, C, H, W, K,
                              float(down_ratio), flags, _lib.ptr(boxes), _lib.ptr(scores),
                              _lib.ptr(classes), _lib.ptr(inds), _lib.ptr(ws),
                              ctypes.c_size_t(ws.numel()), _lib.stream_ptr(dev))
    _lib.check(rc, "ctdet_decode")
    return boxes, scores, classes, inds


def head_decode_fusable(Cf, C, H, W):
    """Shapes `ctdet_head_decode_fused` takes (include/ctdet_b200.h): a map row is one tensor-core tile of <= 128 pixels."""
    return 1 <= W <= 128 and Cf % 64 == 0 and C % 16 == 0 and C <= 128 and \
        1024 + (Cf // 64) * C * 128 + 3 * C * 512 + 2 * 16384 + 8 * C + 5120 + 8192 <= 226 * 1024


def hm_head_decode_batch(feat, weight, bias, wh, reg=None, down_ratio=1, cat_spec_wh=False, K=100, return_logits=False):
    """`ctdet_decode(Conv2d(Cf, C, 1)(feat).sigmoid_(), wh, reg, ...)` (centernet.py:118-123, :152, :161, :426-458) for a
    batch, without the logits in HBM: the 1x1 convolution runs per map row on the tensor cores (bf16 operands) and the
    decode's peak test / threshold sweep on a three-row window in shared memory.  feat f32 [B,Cf,H,W], W <= 128; weight f32 [C,Cf]
    or [C,Cf,1,1]; bias f32 [C] or None.  Returns boxes, scores, classes, inds as ctdet_decode_batch (+ the logits
    f32 [B,C,H,W] with return_logits=True: an extra pass, for verification)."""
    _lib.require_cuda("hm_head_decode", feat, weight, bias, wh, reg)
    lib = _lib.load()
    feat = _lib.f32c(feat, "feat")
    if weight.dim() == 4:
        if weight.shape[2:] != (1, 1):
            raise NotImplementedError("hm_head_decode fuses the head's final 1x1 convolution")
        weight = weight.reshape(weight.shape[0], weight.shape[1])
    weight = _lib.f32c(weight, "weight")
    bias = None if bias is None else _lib.f32c(bias, "bias")
    wh = _lib.f32c(wh, "wh")
    reg = None if reg is None else _lib.f32c(reg, "reg")
    B, Cf, H, W = feat.shape
    C = weight.shape[0]
    if weight.shape[1] != Cf:
        raise RuntimeError("weight must be [%d,%d]" % (C, Cf))
    want_wh = 2 * C if cat_spec_wh else 2
    if tuple(wh.shape) != (B, want_wh, H, W):
        raise RuntimeError("wh must be [%d,%d,%d,%d], got %s" % (B, want_wh, H, W, tuple(wh.shape)))
    if reg is not None and tuple(reg.shape) != (B, 2, H, W):
        raise RuntimeError("reg must be [%d,2,%d,%d]" % (B, H, W))
    if K > H * W:
        raise RuntimeError("selected index k out of range")
    if not head_decode_fusable(Cf, C, H, W):
        raise NotImplementedError("hm_head_decode_batch needs W <= 128, Cf %% 64 == 0, C %% 16 == 0, C <= 128 (got Cf=%d C=%d "
                                  "W=%d): run the convolution and ctdet_decode_batch(logits=True)" % (Cf, C, W))
    dev = feat.device
    boxes = torch.empty((B, K, 4), dtype=torch.float32, device=dev)
    scores = torch.empty((B, K), dtype=torch.float32, device=dev)
    classes = torch.empty((B, K), dtype=torch.int32, device=dev)
    inds = torch.empty((B, K), dtype=torch.int64, device=dev)
    logits = torch.empty((B, C, H, W), dtype=torch.float32, device=dev) if return_logits else None
    nbytes = lib.ctdet_head_decode_fused_workspace_bytes(B, Cf, C, H, W, K)
    ws = _lib.zero_workspace("head_decode%d" % B, max(nbytes, 8), dev)
    with torch.cuda.device(dev):
        rc = lib.ctdet_head_decode_fused(_lib.ptr(feat), _lib.ptr(weight), _lib.ptr(bias), _lib.ptr(wh), _lib.ptr(reg),
                                         B, Cf, C, H, W, K, float(down_ratio), FLAG_CAT_SPEC_WH if cat_spec_wh else 0,
                                         _lib.ptr(boxes), _lib.ptr(scores), _lib.ptr(classes), _lib.ptr(inds),
                                         _lib.ptr(logits), _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(dev))
    _lib.check(rc, "ctdet_head_decode_fused")
    if return_logits:
        return boxes, scores, classes, inds, logits
    return boxes, scores, classes, inds


def ctdet_decode(heat, wh, reg=None, down_ratio=1, cat_spec_wh=False, K=100):
    """Drop-in for centernet.py:426 (batch must be 1, as the reference's `view(K)` at :450-456
    requires): returns (bboxes [K,4], scores [K], clses [K] int32)."""
    if heat.shape[0] != 1:
        raise RuntimeError("shape '[%d]' is invalid for input of size %d" % (K, heat.shape[0] * K))
    boxes, scores, classes, _ = ctdet_decode_batch(heat, wh, reg, down_ratio, cat_spec_wh, K)
    return boxes[0], scores[0], classes[0]


def _nms(heat, kernel=3):
    """Drop-in for centernet.py:399: heat * (max_pool3x3(heat) == heat)."""
    if kernel != 3:
        raise NotImplementedError("only the 3x3 peak NMS of ctdet is implemented")
    _lib.require_cuda("_nms", heat)
    lib = _lib.load()
    heat = _lib.f32c(heat, "heat")
    B, C, H, W = heat.shape
    out = torch.empty_like(heat)
    with torch.cuda.device(heat.device):
        rc = lib.ctdet_nms(_lib.ptr(heat), B, C, H, W, _lib.ptr(out), _lib.stream_ptr(heat.device))
    _lib.check(rc, "ctdet_nms")
    return out


def _topk(scores, K=40):
    """Drop-in for centernet.py:408: (topk_score, topk_inds, topk_clses, topk_ys, topk_xs)."""
    B, C, H, W = scores.shape
    dummy = torch.zeros((B, 2, H, W), dtype=torch.float32, device=scores.device)
    _, s, cls, inds = ctdet_decode_batch(scores, dummy, None, 1, False, K, nms=False)
    ys = (inds // W).to(torch.float32)
    xs = (inds % W).to(torch.float32)
    return s, inds, cls, ys, xs


def _transpose_and_gather_feat(feat, ind):
    """Drop-in for centernet.py:383: feat [B,Ch,H,W], ind [B,K] -> [B,K,Ch] (no NHWC copy)."""
    _lib.require_cuda("_transpose_and_gather_feat", feat, ind)
    lib = _lib.load()
    feat = _lib.f32c(feat, "feat")
    ind = ind.to(torch.int64).contiguous()
    B, Ch = feat.shape[0], feat.shape[1]
    HW = feat.shape[2] * feat.shape[3]
    K = ind.shape[1]
    out = torch.empty((B, K, Ch), dtype=torch.float32, device=feat.device)
    with torch.cuda.device(feat.device):
        rc = lib.ctdet_gather_feat(_lib.ptr(feat), _lib.ptr(ind), B, Ch, HW, K, _lib.ptr(out),
                                   _lib.stream_ptr(feat.device))
    _lib.check(rc, "ctdet_gather_feat")
    return out


def postprocess_batch(boxes, scores, classes, sizes, max_dets=100, score_thresh=0.05):
    """Batched inference_single_image (:251-261) + detector_postprocess
    (modeling/postprocessing.py:10-74), in place.  sizes i32[B,4] = (in_h, in_w, out_h, out_w).
    Returns n_keep i32[B]; survivors are compacted to the front of each row."""
    _lib.require_cuda("postprocess", boxes, scores, classes, sizes)
    lib = _lib.load()
    B, K = scores.shape
    n_keep = torch.empty((B,), dtype=torch.int32, device=scores.device)
    sizes = sizes.to(torch.int32).contiguous()
    with torch.cuda.device(scores.device):
        rc = lib.ctdet_postprocess(_lib.ptr(boxes), _lib.ptr(scores), _lib.ptr(classes),
                                   _lib.ptr(sizes), B, K, int(max_dets), float(score_thresh),
                                   _lib.ptr(n_keep), _lib.stream_ptr(scores.device))
    _lib.check(rc, "ctdet_postprocess")
    return n_keep
