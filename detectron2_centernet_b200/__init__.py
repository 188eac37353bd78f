"""detectron2-centernet_b200: the CenterNet "ctdet" heatmap hot path as hand-written sm_100a CUDA
kernels behind the reference's own operator surface (see DESIGN.md / INTEGRATION.md).

    from detectron2_centernet_b200 import (gen_heatmap, FocalLoss, RegL1Loss, ctdet_decode, _nms, _topk,
                                           _transpose_and_gather_feat, DeformConvV2, modulated_deform_conv)

Importing the package does not load the CUDA library; the first op call does, and raises if
libctdet_b200.so is missing (there is no CPU / PyTorch fallback).
"""
from .target import gen_heatmap, gen_heatmap_batch                                    # noqa: F401
from .losses import FocalLoss, RegL1Loss, _neg_loss, ctdet_losses, ctdet_losses_from_objects                      # noqa: F401
from .decode import (ctdet_decode, ctdet_decode_batch, _nms, _topk, _transpose_and_gather_feat,   # noqa: F401
                     postprocess_batch)
from .deform_conv import DCN, DeformConvV2, ModulatedDeformConv, modulated_deform_conv, dcnv2_forward, dcnv2_backward   # noqa: F401
from .neck import CtdetHead, DLA34Neck, DLAUp, IDAUp, fill_up_weights                 # noqa: F401
from .structures import Boxes, ImageList, Instances                                    # noqa: F401
from .preprocess import preprocess_image                                               # noqa: F401
from .centernet import CenterNetHotPath                                                # noqa: F401
from .config import get_cfg                                                            # noqa: F401
