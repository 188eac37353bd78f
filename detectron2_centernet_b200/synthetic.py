"""Seeded synthetic COCO-shaped inputs for the ctdet hot path.

Host-side numpy only; used by bench.py, the tests and oracle/gen_golden.py so that every leg
(CUDA path, oracle, reference) sees identical bytes.
"""
import numpy as np


def make_objects(rng, input_size, n=None, n_max=40, num_classes=80, degenerate_frac=0.02):
    """Random ground-truth boxes for one image: fp32 XYXY in input pixels clipped to the image
    (so some are border-truncated) and int64 classes.  ~2% are degenerate (zero width or
    height) to exercise the skip at detection_utils.py:621."""
    S = float(input_size)
    if n is None:
        n = int(rng.integers(1, n_max + 1))
    w = rng.uniform(8.0, 0.4 * S, n)
    h = rng.uniform(8.0, 0.4 * S, n)
    cx = rng.uniform(0.0, S, n)
    cy = rng.uniform(0.0, S, n)
    deg = rng.random(n) < degenerate_frac
    w = np.where(deg & (rng.random(n) < 0.5), 0.0, w)
    h = np.where(deg & (w > 0), 0.0, h)
    boxes = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
    boxes = np.clip(boxes, 0.0, S).astype(np.float32)
    classes = rng.integers(0, num_classes, n).astype(np.int64)
    return boxes, classes


def make_batch_objects(seed, B, input_size, n=None, n_max=40, num_classes=80):
    rng = np.random.default_rng(seed)
    bl, cl = [], []
    for _ in range(B):
        b, c = make_objects(rng, input_size, n=n, n_max=n_max, num_classes=num_classes)
        bl.append(b)
        cl.append(c)
    return bl, cl


def pack_objects(boxes_list, classes_list):
    """Ragged -> flat: boxes f32[total,4], classes i64[total], offsets i32[B+1]."""
    counts = [len(c) for c in classes_list]
    offsets = np.zeros(len(counts) + 1, np.int32)
    offsets[1:] = np.cumsum(counts)
    if offsets[-1] == 0:
        return np.zeros((0, 4), np.float32), np.zeros((0,), np.int64), offsets
    return (np.concatenate([np.asarray(b, np.float32).reshape(-1, 4) for b in boxes_list]),
            np.concatenate([np.asarray(c, np.int64).reshape(-1) for c in classes_list]), offsets)


def make_head_outputs(seed, B, C, H, W, boxes_list=None, tie_free=True, down=4):
    """hm logits ~ N(-2.19, 1.5^2) (the -2.19 bias prior, centernet.py:123) with planted peaks
    ~U(0,4) at ground-truth centres; wh ~ U(0,64); reg ~ U(0,1).  tie_free adds a tiny
    index-proportional jitter and keeps logits < 9 (no clamp saturation at 1-1e-4)."""
    rng = np.random.default_rng(seed)
    hm = rng.normal(-2.19, 1.5, (B, C, H, W)).astype(np.float32)
    if boxes_list is not None:
        for b, boxes in enumerate(boxes_list):
            for k in range(len(boxes)):
                x = int(min(max((boxes[k][0] + boxes[k][2]) / (2 * down), 0), W - 1))
                y = int(min(max((boxes[k][1] + boxes[k][3]) / (2 * down), 0), H - 1))
                hm[b, int(rng.integers(0, C)), y, x] = rng.uniform(0.0, 4.0)
    if tie_free:
        n = C * H * W
        jit = (np.arange(n, dtype=np.float64) / n * 1e-3).astype(np.float32).reshape(C, H, W)
        hm = np.minimum(hm + jit[None], np.float32(8.9)).astype(np.float32)
    wh = rng.uniform(0.0, 64.0, (B, 2, H, W)).astype(np.float32)
    reg = rng.uniform(0.0, 1.0, (B, 2, H, W)).astype(np.float32)
    return hm, wh, reg


def sigmoid_clamp(x):
    """clamp(sigmoid(x), 1e-4, 1-1e-4) in fp32 (centernet.py:161), numpy."""
    x = np.asarray(x, np.float32)
    with np.errstate(over="ignore"):
        p = (np.float32(1) / (np.float32(1) + np.exp(-x))).astype(np.float32)
    return np.clip(p, np.float32(1e-4), np.float32(1 - 1e-4)).astype(np.float32)
