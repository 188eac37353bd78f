"""preprocess_image (meta_arch/centernet.py:173-184 + structures/image_list.py:58-130): the oracle
against the committed outputs of the reference's own code, and the CUDA kernel against both.
Tolerance: 0 -- the kernel follows the reference's fp32 operation order with IEEE division."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ctdet_oracle as O                      # noqa: E402
from oracle import gen_golden_preprocess as GP            # noqa: E402
from oracle import ref_loader as RL                       # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "preprocess_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("case", sorted(GP.CASES))
def test_oracle_matches_reference_outputs(case, golden):
    imgs, div = GP.build_images(case)
    t, sizes = O.preprocess_image(imgs, GP.MEAN, GP.STD, div)
    assert np.array_equal(t.numpy(), golden[case + "/tensor"])
    assert np.array_equal(np.asarray(sizes, np.int32), golden[case + "/sizes"])


@pytest.mark.skipif(not RL.available(), reason="reference tree not present (GPU box)")
def test_oracle_vs_live_reference():
    g = torch.Generator().manual_seed(7)
    imgs = [torch.randint(0, 256, (3, int(h), int(w)), generator=g, dtype=torch.uint8)
            for h, w in torch.randint(1, 70, (5, 2), generator=g).tolist()]
    for div in (0, 16, 32):
        a, sa = RL.reference_preprocess(imgs, GP.MEAN, GP.STD, div)
        b, sb = O.preprocess_image(imgs, GP.MEAN, GP.STD, div)
        assert torch.equal(a, b) and sa == sb


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(GP.CASES))
def test_kernel_matches_reference_outputs(case, golden):
    from detectron2_centernet_b200 import preprocess_image
    imgs, div = GP.build_images(case)
    il = preprocess_image([t.cuda() for t in imgs], GP.MEAN, GP.STD, div)
    assert np.array_equal(il.tensor.cpu().numpy(), golden[case + "/tensor"])
    assert np.array_equal(np.asarray(il.image_sizes, np.int32), golden[case + "/sizes"])
    h, w = il.image_sizes[0]
    assert il[0].shape[-2:] == (h, w)


@pytest.mark.gpu
def test_kernel_full_size_and_errors():
    from detectron2_centernet_b200 import CenterNetHotPath, preprocess_image
    g = torch.Generator().manual_seed(3)
    sizes = [(512, 512), (480, 511), (333, 500), (512, 384)]
    imgs = [torch.randint(0, 256, (3, h, w), generator=g, dtype=torch.uint8) for h, w in sizes]
    hot = CenterNetHotPath()
    il = hot.preprocess_image([{"image": t} for t in imgs], device="cuda")
    ref, rs = O.preprocess_image(imgs, hot.pixel_mean, hot.pixel_std, hot.size_divisibility)
    assert il.tensor.shape == ref.shape == (4, 3, 512, 512) and il.image_sizes == rs
    assert torch.equal(il.tensor.cpu(), ref)
    # torch's CUDA kernel for `tensor / 255.` multiplies by the rounded reciprocal (<= 1 ulp away from
    # the IEEE quotient the CPU path and this kernel compute): stated tolerance 1e-6
    dev_ref = [((t.cuda() / 255. - torch.tensor(hot.pixel_mean, device="cuda").view(-1, 1, 1))
                / torch.tensor(hot.pixel_std, device="cuda").view(-1, 1, 1)) for t in imgs]
    for i, (h, w) in enumerate(sizes):
        torch.testing.assert_close(il.tensor[i, :, :h, :w], dev_ref[i], rtol=1e-6, atol=1e-6)
    with pytest.raises(NotImplementedError):
        preprocess_image(imgs, GP.MEAN, GP.STD, 16)            # CPU tensors: no fallback
    with pytest.raises(ValueError):
        preprocess_image([], GP.MEAN, GP.STD, 16)
    with pytest.raises(RuntimeError):
        preprocess_image([imgs[0].cuda(), imgs[1].cuda().float()], GP.MEAN, GP.STD, 16)
