"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Loads the *reference's own* ctdet hot-path functions from the read-only tree at
/root/reference without importing the detectron2 package (which needs fvcore, yacs,
pycocotools and a built detectron2._C, none of which exist in this image).

The named FunctionDef / ClassDef nodes are AST-extracted from

  detectron2/data/detection_utils.py        gen_heatmap :600, gaussian_radius :654,
                                            gaussian2D :682, draw_umich_gaussian :690
  detectron2/modeling/meta_arch/centernet.py  FocalLoss :323, _neg_loss :333, RegL1Loss :372,
                                            _transpose_and_gather_feat :383, _gather_feat :389,
                                            _nms :399, _topk :408, ctdet_decode :426

and exec'd unmodified with {torch, nn, F, np} in scope.  `_neg_loss` calls `.cuda()` on its
alpha tensor (:342-349); on a CPU-only host `Tensor.cuda` is shimmed to identity while the
reference function runs.

This only works where /root/reference exists (the build container).  It is used by
oracle/gen_golden.py to produce tests/golden/*.npz and by the CPU tests (when the tree is
present) to pin oracle/ctdet_oracle.py against the real reference.  Nothing on the GPU box
may depend on it.
"""
import ast
import contextlib
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

REF_ROOT = os.environ.get("CTDET_REFERENCE_ROOT", "/root/reference")

_WANTED = {
    "detectron2/data/detection_utils.py": [
        "gen_heatmap", "gaussian_radius", "gaussian2D", "draw_umich_gaussian"],
    "detectron2/modeling/meta_arch/centernet.py": [
        "FocalLoss", "_neg_loss", "RegL1Loss", "_transpose_and_gather_feat",
        "_gather_feat", "_nms", "_topk", "ctdet_decode"],
}


def available():
    return all(os.path.isfile(os.path.join(REF_ROOT, p)) for p in _WANTED)


def _extract(path, names):
    with open(os.path.join(REF_ROOT, path), "r") as f:
        src = f.read()
    tree = ast.parse(src)
    picked = [n for n in tree.body
              if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names]
    missing = set(names) - {n.name for n in picked}
    if missing:
        raise RuntimeError("reference symbols not found in %s: %s" % (path, sorted(missing)))
    return ast.Module(body=picked, type_ignores=[])


_CACHE = {}


def load():
    """Returns a dict name -> reference object (functions / nn.Module classes)."""
    if "gen_heatmap" in _CACHE:
        return _CACHE
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    ns = {"torch": torch, "nn": nn, "F": F, "np": np, "__name__": "ctdet_reference"}
    for path, names in _WANTED.items():
        mod = _extract(path, names)
        exec(compile(mod, os.path.join(REF_ROOT, path), "exec"), ns)
    for names in _WANTED.values():
        for n in names:
            _CACHE[n] = ns[n]
    return _CACHE


def load_image_list():
    """The reference's own ImageList class (detectron2/structures/image_list.py imports nothing but
    torch, so the whole file is exec'd unmodified)."""
    if "ImageList" not in _CACHE:
        path = os.path.join(REF_ROOT, "detectron2/structures/image_list.py")
        ns = {"__name__": "ctdet_reference_image_list"}
        with open(path, "r") as f:
            exec(compile(f.read(), path, "exec"), ns)
        _CACHE["ImageList"] = ns["ImageList"]
    return _CACHE["ImageList"]


def reference_preprocess(images, pixel_mean, pixel_std, size_divisibility):
    """preprocess_image of the reference, centernet.py:177-184, with its own ImageList."""
    ImageList = load_image_list()
    mean = torch.Tensor(list(pixel_mean)).view(-1, 1, 1)                 # :64
    std = torch.Tensor(list(pixel_std)).view(-1, 1, 1)                   # :65
    xs = [x / 255. for x in images]                                      # :177
    xs = [(x - mean) / std for x in xs]                                  # :183
    il = ImageList.from_tensors(xs, size_divisibility)                   # :184
    return il.tensor, [tuple(int(v) for v in s) for s in il.image_sizes]


@contextlib.contextmanager
def cpu_cuda_shim():
    """`_neg_loss` does `torch.tensor(alpha).cuda()`; make that a no-op on CPU-only hosts."""
    if torch.cuda.is_available():
        yield
        return
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        yield
    finally:
        torch.Tensor.cuda = orig


class _Boxes:
    def __init__(self, t):
        self.tensor = t


class DuckInstances:
    """What gen_heatmap reads from an Instances: .gt_boxes.tensor [n,4] f32, .gt_classes [n] i64."""

    def __init__(self, boxes, classes):
        self.gt_boxes = _Boxes(torch.as_tensor(boxes, dtype=torch.float32))
        self.gt_classes = torch.as_tensor(classes, dtype=torch.int64)


class DuckMeta:
    def __init__(self, num_classes):
        self.thing_classes = ["c%d" % i for i in range(num_classes)]


def ref_gen_heatmap(boxes, classes, H, W, C):
    """Runs the reference gen_heatmap (detection_utils.py:600) on CPU tensors."""
    fn = load()["gen_heatmap"]
    return fn(DuckInstances(boxes, classes), np.array([H, W]), DuckMeta(C))


def ref_losses(hm_logits, wh, reg, gt, alpha=(1.0,), weights=(1.0, 0.1, 1.0)):
    """Reference CenterNet.losses (centernet.py:191-212) on given head outputs; returns
    (dict of losses, grads wrt hm_logits/wh/reg)."""
    R = load()
    hm = hm_logits.clone().requires_grad_(True)
    w_ = wh.clone().requires_grad_(True)
    r_ = reg.clone().requires_grad_(True)
    with cpu_cuda_shim():
        crit = R["FocalLoss"](list(alpha))
        regc = R["RegL1Loss"]()
        # `outputs['hm'].sigmoid_()` is in-place on the head output (a non-leaf); mirror with x*1
        hm_loss = crit(torch.clamp((hm * 1).sigmoid_(), min=1e-4, max=1 - 1e-4), gt["hm"])
        wh_loss = regc(w_, gt["reg_mask"], gt["ind"], gt["wh"])
        off_loss = regc(r_, gt["reg_mask"], gt["ind"], gt["reg"])
        out = {"hm_loss": hm_loss * weights[0], "wh_loss": wh_loss * weights[1],
               "off_loss": off_loss * weights[2]}
        (out["hm_loss"] + out["wh_loss"] + out["off_loss"]).backward()
    return out, (hm.grad, w_.grad, r_.grad)


def ref_decode(heat, wh, reg, down_ratio=4, K=100):
    """Reference ctdet_decode (centernet.py:426) looped per image (it is batch=1 only)."""
    fn = load()["ctdet_decode"]
    outs = [fn(heat[b:b + 1], wh[b:b + 1], reg=None if reg is None else reg[b:b + 1],
               down_ratio=down_ratio, K=K) for b in range(heat.shape[0])]
    return (torch.stack([o[0] for o in outs]), torch.stack([o[1] for o in outs]),
            torch.stack([o[2] for o in outs]))


# --------------------------------------------------------------------------------------
# Inference glue: CenterNet.inference / inference_single_image (centernet.py:214-266) and
# detector_postprocess (modeling/postprocessing.py:11-74) with the reference's own Instances / Boxes.
# --------------------------------------------------------------------------------------
def load_structures():
    """The reference's own `Boxes` and `Instances` (detectron2/structures/{boxes,instances}.py import only
    torch / numpy / typing, so both files are exec'd whole and unmodified)."""
    if "Instances" not in _CACHE:
        for fname, names in (("boxes.py", ("Boxes",)), ("instances.py", ("Instances",))):
            path = os.path.join(REF_ROOT, "detectron2/structures", fname)
            ns = {"__name__": "ctdet_reference_" + fname[:-3]}
            with open(path, "r") as f:
                exec(compile(f.read(), path, "exec"), ns)
            for n in names:
                _CACHE[n] = ns[n]
    return _CACHE["Boxes"], _CACHE["Instances"]


def _extract_methods(path, cls_name, names):
    with open(os.path.join(REF_ROOT, path), "r") as f:
        tree = ast.parse(f.read())
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls_name]
    if not cls:
        raise RuntimeError("class %s not found in %s" % (cls_name, path))
    picked = [n for n in cls[0].body if isinstance(n, ast.FunctionDef) and n.name in names]
    missing = set(names) - {n.name for n in picked}
    if missing:
        raise RuntimeError("reference methods not found in %s.%s: %s" % (path, cls_name, sorted(missing)))
    return ast.Module(body=picked, type_ignores=[])


def load_inference():
    """-> (inference, inference_single_image, detector_postprocess): the reference's unmodified FunctionDef nodes.
    The two methods are plain functions taking `self` first (call them with `DummyCenterNet`)."""
    if "inference_single_image" not in _CACHE:
        R = load()
        Boxes, Instances = load_structures()
        path = "detectron2/modeling/meta_arch/centernet.py"
        ns = {"torch": torch, "nn": nn, "F": F, "np": np, "Instances": Instances, "Boxes": Boxes,
              "ctdet_decode": R["ctdet_decode"], "__name__": "ctdet_reference_inference"}
        exec(compile(_extract_methods(path, "CenterNet", ["inference", "inference_single_image"]),
                     os.path.join(REF_ROOT, path), "exec"), ns)
        _CACHE["inference"], _CACHE["inference_single_image"] = ns["inference"], ns["inference_single_image"]
        path = "detectron2/modeling/postprocessing.py"
        ns2 = {"torch": torch, "F": F, "Instances": Instances, "__name__": "ctdet_reference_postprocessing",
               # only reached for pred_masks, which ctdet never produces (postprocessing.py:60-66)
               "paste_masks_in_image": None, "retry_if_cuda_oom": lambda f: f}
        exec(compile(_extract(path, ["detector_postprocess"]), os.path.join(REF_ROOT, path), "exec"), ns2)
        _CACHE["detector_postprocess"] = ns2["detector_postprocess"]
    return _CACHE["inference"], _CACHE["inference_single_image"], _CACHE["detector_postprocess"]


class _Backbone:
    def __init__(self, down_ratio):
        self.down_ratio = down_ratio


class DummyCenterNet:
    """The four attributes `inference_single_image` reads from `self` (centernet.py:245-258), set from the
    config fields the reference's __init__ reads (:50-57)."""

    def __init__(self, down_ratio=4, topk_candidates=100, max_detections_per_image=100, score_threshold=0.05):
        self.backbone = _Backbone(down_ratio)
        self.topk_candidates = topk_candidates
        self.max_detections_per_image = max_detections_per_image
        self.score_threshold = score_threshold

    def inference_single_image(self, output, image_size):
        return load_inference()[1](self, output, image_size)


def ref_inference(hm_logits, wh, reg, image_sizes, output_sizes=None, **cfg):
    """The reference's eval path from the head outputs on: sigmoid_ + clamp (centernet.py:161), `inference`
    (:214-234) and, when output_sizes is given, `detector_postprocess` per image (:169-171 of the reference's
    forward).  Returns a list of the reference's own Instances."""
    inference, _, post = load_inference()
    net = DummyCenterNet(**cfg)
    out = {"hm": torch.clamp(hm_logits.clone().sigmoid_(), min=1e-4, max=1 - 1e-4), "wh": wh, "reg": reg}
    res = inference(net, out, image_sizes)
    if output_sizes is not None:
        res = [post(r, int(o[0]), int(o[1])) for r, o in zip(res, output_sizes)]
    return res
