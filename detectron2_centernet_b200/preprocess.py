"""CenterNet.preprocess_image on the GPU (detectron2/modeling/meta_arch/centernet.py:173-189):
`x / 255`, `(x - pixel_mean) / pixel_std`, `ImageList.from_tensors(images, size_divisibility)`
(structures/image_list.py:58-130) for the whole batch in one launch of `ctdet_preprocess`.
"""
import torch

from . import _lib
from .structures import ImageList

_IMAGE_U8, _IMAGE_F32 = 0, 1


def preprocess_image(images, pixel_mean, pixel_std, size_divisibility=0):
    """images: list of CUDA tensors [C,h_i,w_i], all uint8 (what the dataset mapper yields) or all
    float32; pixel_mean / pixel_std: per-channel sequences or tensors (cfg.MODEL.PIXEL_MEAN / _STD).
    Returns the ImageList of the reference: tensor f32 [N,C,Hp,Wp] zero-padded bottom/right to the
    batch maximum rounded up to `size_divisibility`, and the original sizes."""
    if not isinstance(images, (list, tuple)) or len(images) == 0:
        raise ValueError("preprocess_image: a non-empty list of image tensors is required")   # image_list.py:77-78
    _lib.require_cuda("preprocess_image", *images)
    dev = images[0].device
    C = images[0].shape[0]
    dt = images[0].dtype
    if dt not in (torch.uint8, torch.float32):
        raise RuntimeError("preprocess_image: images must be uint8 or float32, got %s" % dt)
    imgs = []
    for t in images:
        if t.dim() != 3 or t.shape[0] != C or t.dtype != dt:
            raise RuntimeError("preprocess_image: every image must be [C,h,w] of one dtype (image_list.py:79-81)")
        imgs.append(t.contiguous())
    sizes = [(int(t.shape[1]), int(t.shape[2])) for t in imgs]
    Hp, Wp = max(s[0] for s in sizes), max(s[1] for s in sizes)
    if size_divisibility > 1:                                          # image_list.py:99-102
        d = int(size_divisibility)
        Hp, Wp = (Hp + d - 1) // d * d, (Wp + d - 1) // d * d
    if Wp % 4 != 0:
        raise NotImplementedError("preprocess_image: padded width must be a multiple of 4 (size_divisibility 16/32)")
    mean = torch.as_tensor(pixel_mean, dtype=torch.float32).reshape(-1).to(dev)
    std = torch.as_tensor(pixel_std, dtype=torch.float32).reshape(-1).to(dev)
    if mean.numel() != C or std.numel() != C:
        raise RuntimeError("preprocess_image: pixel_mean / pixel_std need one value per channel")
    # one small H2D copy: the pointer table and the sizes
    # (little-endian: the int64 h | w << 32 is the int32 pair {h, w} the kernel reads)
    table = torch.tensor([t.data_ptr() for t in imgs] + [h | (w << 32) for h, w in sizes], dtype=torch.int64)
    table = table.to(dev, non_blocking=True)
    ptrs = table[:len(imgs)]
    sizes_dev = table[len(imgs):].view(torch.int32)
    out = torch.empty((len(imgs), C, Hp, Wp), dtype=torch.float32, device=dev)
    lib = _lib.load()
    with torch.cuda.device(dev):
        _lib.check(lib.ctdet_preprocess(_lib.ptr(ptrs), _lib.ptr(sizes_dev), _IMAGE_U8 if dt == torch.uint8 else _IMAGE_F32,
                                        len(imgs), C, Hp, Wp, _lib.ptr(mean), _lib.ptr(std), _lib.ptr(out),
                                        _lib.stream_ptr(dev)), "ctdet_preprocess")
    return ImageList(out, sizes)
