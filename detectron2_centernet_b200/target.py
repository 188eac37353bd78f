"""Gaussian target generation on the GPU -- the operator surface of
detectron2/data/detection_utils.py:600 (`gen_heatmap`) and the batching of
detectron2/modeling/meta_arch/centernet.py:197-201, backed by `ctdet_target_gen`.
"""
import numpy as np
import torch

from . import _lib

MAX_OBJS = 128   # detection_utils.py:611-616


def target_gen(boxes, classes, offsets, B, C, H, W, out=None):
    """Low-level call.  boxes f32[total,4] (XYXY input pixels), classes i64[total],
    offsets i32[B+1], all CUDA.  Returns dict hm/reg_mask/ind/wh/reg (batched) + num_pos i32[B]."""
    _lib.require_cuda("ctdet target generation", boxes, classes, offsets)
    lib = _lib.load()
    dev = offsets.device
    if boxes.dtype != torch.float32 or classes.dtype != torch.int64 or offsets.dtype != torch.int32:
        raise RuntimeError("target_gen expects boxes float32, classes int64, offsets int32")
    if offsets.numel() != B + 1:
        raise RuntimeError("offsets must have B+1 entries")
    boxes = boxes.contiguous()
    classes = classes.contiguous()
    offsets = offsets.contiguous()
    if out is None:
        out = {
            "hm": torch.empty((B, C, H, W), dtype=torch.float32, device=dev),
            "reg_mask": torch.empty((B, MAX_OBJS), dtype=torch.uint8, device=dev),
            "ind": torch.empty((B, MAX_OBJS), dtype=torch.int64, device=dev),
            "wh": torch.empty((B, MAX_OBJS, 2), dtype=torch.float32, device=dev),
            "reg": torch.empty((B, MAX_OBJS, 2), dtype=torch.float32, device=dev),
            "num_pos": torch.empty((B,), dtype=torch.int32, device=dev),
        }
    with torch.cuda.device(dev):
        rc = lib.ctdet_target_gen(
            _lib.ptr(boxes), _lib.ptr(classes), _lib.ptr(offsets), B, C, H, W,
            _lib.ptr(out["hm"]), _lib.ptr(out["reg_mask"]), _lib.ptr(out["ind"]),
            _lib.ptr(out["wh"]), _lib.ptr(out["reg"]), _lib.ptr(out["num_pos"]),
            _lib.stream_ptr(dev))
    _lib.check(rc, "ctdet_target_gen")
    return out


def _num_classes(meta):
    return len(meta.thing_classes) if hasattr(meta, "thing_classes") else int(meta)


def pack_instances(instances, device=None):
    """list of per-image Instances (fields gt_boxes.tensor, gt_classes) -> CSR device tensors
    (boxes f32 [total,4], classes i64 [total], offsets i32 [B+1]) as the C ABI takes them."""
    B = len(instances)
    boxes = [x.gt_boxes.tensor for x in instances]
    classes = [x.gt_classes for x in instances]
    if device is None:
        device = boxes[0].device if B else torch.device("cuda")
    device = torch.device(device)
    if device.type != "cuda":
        raise NotImplementedError("gen_heatmap is not supported on CPU tensors (CUDA sm_100a only)")
    counts = [int(c.shape[0]) for c in classes]
    offsets = torch.from_numpy(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)).to(
        device, non_blocking=True)
    if sum(counts) == 0:
        fb = torch.zeros((1, 4), dtype=torch.float32, device=device)
        fc = torch.zeros((1,), dtype=torch.int64, device=device)
    else:
        fb = torch.cat([b.reshape(-1, 4) for b in boxes]).to(device=device, dtype=torch.float32)
        fc = torch.cat([c.reshape(-1) for c in classes]).to(device=device, dtype=torch.int64)
    return fb, fc, offsets


def gen_heatmap_batch(instances, output_shape, meta, device=None):
    """Batched gen_heatmap: `instances` is a list of per-image Instances (fields gt_boxes.tensor,
    gt_classes).  One launch for the whole batch; returns the five tensors already concatenated as
    CenterNet.losses builds them (centernet.py:197-201), plus num_pos i32[B]."""
    H, W = int(output_shape[0]), int(output_shape[1])
    fb, fc, offsets = pack_instances(instances, device)
    return target_gen(fb, fc, offsets, len(instances), _num_classes(meta), H, W)


def gen_heatmap(instances, output_shape, meta):
    """Drop-in for detection_utils.py:600: one image -> dict hm[C,H,W], reg_mask[128] u8,
    ind[128] i64, wh[128,2], reg[128,2] (CUDA tensors)."""
    out = gen_heatmap_batch([instances], output_shape, meta)
    return {k: out[k][0] for k in ("hm", "reg_mask", "ind", "wh", "reg")}
