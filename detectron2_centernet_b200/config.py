"""The config keys the ctdet hot path honours (`MODEL.CENTERNET.*`, `TEST.DETECTIONS_PER_IMAGE`),
with the reference's defaults (detectron2/config/defaults.py:463-494, :630) and `_BASE_` inheritance
as in projects/CenterNet/configs/COCO-Detection/ctdet_dla_34_1x.yaml.  yacs is not needed: plain
PyYAML + a defaults table.  Python tuples in the YAML files ("(a, b)") are kept as strings; the hot
path never reads those keys."""
import copy
import os

import yaml

DEFAULTS = {
    "MODEL": {
        "META_ARCHITECTURE": "CenterNet",
        "PIXEL_MEAN": [0.408, 0.447, 0.470],
        "PIXEL_STD": [0.289, 0.274, 0.278],
        "CENTERNET": {
            "NUM_CLASSES": 80,
            "DOWN_RATIO": 4,
            "HEAD_CONV": 256,
            "FINAL_KERNEL": 1,
            "SIZE_DIVISIBILITY": 32,
            "HM_WEIGHT": 1,
            "WH_WEIGHT": 0.1,
            "OFF_WEIGHT": 1,
            "FOCAL_LOSS_ALPHA": [0.25],
            "TASK": {"HM": 80, "WH": 2, "REG": 2},
            "SCORE_THRESH_TEST": 0.05,
            "TOPK_CANDIDATES_TEST": 100,
        },
    },
    "TEST": {"DETECTIONS_PER_IMAGE": 100},
}


def _merge(dst, src):
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dst.get(k), dict):
            _merge(dst[k], v)
        else:
            dst[k] = copy.deepcopy(v)
    return dst


def _load_yaml(path):
    with open(path) as f:
        cfg = yaml.safe_load(f) or {}
    base = cfg.pop("_BASE_", None)
    if base:
        if not os.path.isabs(base):
            base = os.path.join(os.path.dirname(path), base)
        cfg = _merge(_load_yaml(base), cfg)
    return cfg


def get_cfg(path=None, overrides=None):
    """Defaults, then the YAML file (with `_BASE_`), then a nested dict of overrides."""
    cfg = copy.deepcopy(DEFAULTS)
    if path:
        _merge(cfg, _load_yaml(path))
    if overrides:
        _merge(cfg, overrides)
    return cfg
