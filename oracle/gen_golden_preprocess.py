"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/preprocess_golden.npz by running the reference's
own preprocess_image arithmetic (meta_arch/centernet.py:177-184) with its own ImageList.from_tensors
(structures/image_list.py, exec'd unmodified by oracle/ref_loader.py) on seeded images.

Run in the build container (needs /root/reference):   python -m oracle.gen_golden_preprocess
"""
import os

import numpy as np
import torch

from . import ref_loader as RL

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "preprocess_golden.npz")
MEAN, STD = [0.408, 0.447, 0.470], [0.289, 0.274, 0.278]      # projects/CenterNet/configs/COCO-Detection/Base-CenterNet.yaml:5-6
CASES = {                                                    # name -> (sizes, dtype, size_divisibility)
    "u8_ragged": ([(21, 30), (32, 17), (9, 9)], "u8", 16),
    "u8_single": ([(24, 40)], "u8", 32),
    "f32_ragged": ([(5, 12), (16, 16)], "f32", 16),
}


def build_images(name):
    sizes, dtype, div = CASES[name]
    g = torch.Generator().manual_seed(sum(ord(c) for c in name))
    if dtype == "u8":
        return [torch.randint(0, 256, (3, h, w), generator=g, dtype=torch.uint8) for h, w in sizes], div
    return [torch.rand((3, h, w), generator=g) * 255.0 for h, w in sizes], div


def main():
    out = {}
    for name in CASES:
        imgs, div = build_images(name)
        t, sizes = RL.reference_preprocess(imgs, MEAN, STD, div)
        out[name + "/tensor"] = t.numpy()
        out[name + "/sizes"] = np.asarray(sizes, np.int32)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
