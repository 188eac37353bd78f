"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

fp32 stand-ins for the missing `dcn_v2.DCN` submodule and for `DeformConvV2`
(detectron2/layers/deform_conv.py:498-519), built on torchvision.ops.deform_conv2d(mask=...) -- the
implementation oracle/dcn_oracle.c is pinned against -- with the reference's parameter names, so that
(a) the reference's own IDAUp / DLAUp classes (AST-extracted from detectron2/modeling/backbone/dla.py
in the build container) can be instantiated and compared with the mirror in
detectron2_centernet_b200/neck.py, and (b) the B200 DeformConvV2 can be checked inside the whole neck.
"""
import ast
import math
import os

import numpy as np
import torch
from torch import nn
from torchvision.ops import deform_conv2d

from . import ref_loader as RL


class OracleDCN(nn.Module):
    """dcn_v2.DCN (call site deform_conv.py:505-513): 27-channel offset/mask conv (zero-initialised),
    offset = cat(o1, o2), mask = sigmoid (convention of modeling/backbone/resnet.py:308-313)."""

    def __init__(self, chi, cho, kernel_size=(3, 3), stride=1, padding=1, dilation=1, deformable_groups=1):
        super().__init__()
        self.stride, self.padding, self.dilation = stride, padding, dilation
        self.weight = nn.Parameter(torch.empty(cho, chi, *kernel_size))
        self.bias = nn.Parameter(torch.zeros(cho))
        self.conv_offset_mask = nn.Conv2d(chi, deformable_groups * 3 * kernel_size[0] * kernel_size[1],
                                          kernel_size=kernel_size, stride=stride, padding=padding, bias=True)
        stdv = 1.0 / math.sqrt(chi * kernel_size[0] * kernel_size[1])
        self.weight.data.uniform_(-stdv, stdv)
        self.conv_offset_mask.weight.data.zero_()
        self.conv_offset_mask.bias.data.zero_()

    def forward(self, x):
        o1, o2, mask = torch.chunk(self.conv_offset_mask(x), 3, dim=1)
        return deform_conv2d(x, torch.cat((o1, o2), dim=1), self.weight, self.bias, stride=self.stride,
                             padding=self.padding, dilation=self.dilation, mask=torch.sigmoid(mask))


class OracleDeformConvV2(nn.Module):
    """deform_conv.py:498-519: `actf` = BatchNorm(momentum 0.1) + ReLU after `conv` = DCN."""

    def __init__(self, chi, cho):
        super().__init__()
        self.actf = nn.Sequential(nn.BatchNorm2d(cho, momentum=0.1), nn.ReLU(inplace=True))
        self.conv = OracleDCN(chi, cho)

    def forward(self, x):
        return self.actf(self.conv(x))


def load_reference_neck():
    """The reference's own fill_up_weights / IDAUp / DLAUp (dla.py:33, :152, :180), exec'd unmodified
    with OracleDeformConvV2 standing in for `detectron2.layers.DeformConvV2`.  Build container only."""
    path = os.path.join(RL.REF_ROOT, "detectron2/modeling/backbone/dla.py")
    with open(path, "r") as f:
        tree = ast.parse(f.read())
    want = {"fill_up_weights", "IDAUp", "DLAUp"}
    picked = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in want]
    assert {n.name for n in picked} == want
    ns = {"torch": torch, "nn": nn, "np": np, "math": math, "DeformConvV2": OracleDeformConvV2,
          "__name__": "ctdet_reference_dla"}
    exec(compile(ast.Module(body=picked, type_ignores=[]), path, "exec"), ns)
    return ns


def randomize(module, seed):
    """Non-trivial weights everywhere (the DCN offset convs and BN statistics start at zero / identity)."""
    g = torch.Generator().manual_seed(seed)
    for name, p in module.named_parameters():
        if "up_" in name:                       # the fixed bilinear kernels stay
            continue
        if name.endswith("conv_offset_mask.weight"):
            p.data.copy_(torch.randn(p.shape, generator=g) * 0.05)
        elif name.endswith("conv_offset_mask.bias"):
            p.data.copy_(torch.randn(p.shape, generator=g) * 0.3)
        elif p.dim() > 1:
            p.data.copy_((torch.rand(p.shape, generator=g) - 0.5) * 2 / math.sqrt(p[0].numel()))
        else:
            p.data.copy_(torch.rand(p.shape, generator=g) * 0.5 + 0.75 if name.endswith("actf.0.weight")
                         else torch.randn(p.shape, generator=g) * 0.1)
    for name, b in module.named_buffers():
        if name.endswith("running_mean"):
            b.copy_(torch.randn(b.shape, generator=g) * 0.1)
        elif name.endswith("running_var"):
            b.copy_(torch.rand(b.shape, generator=g) * 0.5 + 0.75)


def upsample_add_ref(x, weight, skip, f):
    """The two torch ops IDAUp.forward runs (dla.py:175-177) on CPU fp32: `up(x)` with
    up = ConvTranspose2d(o, o, 2f, stride=f, padding=f//2, groups=o, bias=False) (dla.py:166-168), then `+ skip`.
    TEST INFRASTRUCTURE: the checker of ctdet_upsample_add."""
    import torch.nn.functional as F
    y = F.conv_transpose2d(torch.as_tensor(x), torch.as_tensor(weight), None, stride=f, padding=f // 2,
                           groups=weight.shape[0])
    return y if skip is None else y + torch.as_tensor(skip)
