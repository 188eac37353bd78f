"""DCNv2 on the GPU -- the operator surface of detectron2/layers/deform_conv.py:
`modulated_deform_conv` (:306, argument order of `_ModulatedDeformConv.forward` :182-194),
`ModulatedDeformConv` (:406), the missing-submodule `DCN` that `DeformConvV2` instantiates
(:505-513; offset/mask convention as shown in-tree at modeling/backbone/resnet.py:308-313) and
`DeformConvV2(chi, cho)` (:498-519), backed by the tcgen05 implicit-GEMM kernel `dcnv2_forward`.

Parameter names match the reference so state_dicts stay loadable: `conv.weight`, `conv.bias`,
`conv.conv_offset_mask.{weight,bias}`, `actf.0.*`.
The backward (`dcnv2_backward`) runs on the tensor cores as well.
"""
import ctypes
import math

import torch
import torch.nn as nn
from torch.nn.modules.utils import _pair

from . import _lib

FLAG_RELU, FLAG_FP32_BLEND, FLAG_PACKED_WEIGHT = 1, 2, 4

_ws = {}


def _workspace(nbytes, device):
    k = (str(device), torch.cuda.current_stream(device).cuda_stream)
    buf = _ws.get(k)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _ws[k] = buf
    return buf


def dcnv2_pack_weight(weight):
    """The bf16, swizzled tensor-core image of `weight` [Cout,Cin,3,3] that `dcnv2_forward` / `dcnv2_forward_raw27`
    otherwise rebuild on every call (C ABI dcnv2_pack_weight).  Keep it per module and weight version and pass it as
    `packed_weight=`; it is only valid for the weight values it was made from."""
    if not weight.is_cuda:
        raise NotImplementedError("Deformable Conv is not supported on CPUs!")
    lib = _lib.load()
    weight = _lib.f32c(weight.detach(), "weight")
    Cout, Cin, kh, kw = weight.shape
    if (kh, kw) != (3, 3):
        raise NotImplementedError("dcnv2_pack_weight: 3x3 kernels")
    packed = torch.empty(lib.dcnv2_packed_weight_bytes(Cin, Cout), dtype=torch.uint8, device=weight.device)
    with torch.cuda.device(weight.device):
        rc = lib.dcnv2_pack_weight(_lib.ptr(weight), Cin, Cout, _lib.ptr(packed), _lib.stream_ptr(weight.device))
    _lib.check(rc, "dcnv2_pack_weight")
    return packed


def dcnv2_forward(input, offset, mask, weight, bias=None, stride=1, padding=0, dilation=1, groups=1,
                  deformable_groups=1, scale=None, shift=None, relu=False, fp32_blend=False, packed_weight=None):
    """Low-level call.  Optional fused epilogue: (conv + bias) * scale[c] + shift[c], then ReLU.
    packed_weight: `dcnv2_pack_weight(weight)` of the same weights (skips the per-call weight pack)."""
    if not input.is_cuda:
        raise NotImplementedError("Deformable Conv is not supported on CPUs!")    # deform_conv.py:203-204
    lib = _lib.load()
    input = _lib.f32c(input, "input")
    weight = _lib.f32c(weight, "weight")
    offset = _lib.f32c(offset, "offset")
    mask = _lib.f32c(mask, "mask")
    bias = None if bias is None else _lib.f32c(bias, "bias")
    scale = None if scale is None else _lib.f32c(scale, "scale")
    shift = None if shift is None else _lib.f32c(shift, "shift")
    s, p, d = _pair(stride), _pair(padding), _pair(dilation)
    B, Cin, H, W = input.shape
    Cout, cin_g, kh, kw = weight.shape
    if cin_g * groups != Cin:
        raise RuntimeError("weight shape %s does not match input channels %d / groups %d" % (tuple(weight.shape), Cin, groups))
    Ho = (H + 2 * p[0] - (d[0] * (kh - 1) + 1)) // s[0] + 1
    Wo = (W + 2 * p[1] - (d[1] * (kw - 1) + 1)) // s[1] + 1
    if tuple(offset.shape) != (B, 2 * kh * kw * deformable_groups, Ho, Wo):
        raise RuntimeError("offset must be [%d,%d,%d,%d], got %s" % (B, 2 * kh * kw * deformable_groups, Ho, Wo, tuple(offset.shape)))
    if tuple(mask.shape) != (B, kh * kw * deformable_groups, Ho, Wo):
        raise RuntimeError("mask must be [%d,%d,%d,%d], got %s" % (B, kh * kw * deformable_groups, Ho, Wo, tuple(mask.shape)))
    out = torch.empty((B, Cout, Ho, Wo), dtype=torch.float32, device=input.device)
    if out.numel() == 0:
        return out
    nbytes = lib.dcnv2_workspace_bytes(B, Cin, H, W, Cout)
    ws = _workspace(max(nbytes, 16), input.device)
    flags = (FLAG_RELU if relu else 0) | (FLAG_FP32_BLEND if fp32_blend else 0)
    if packed_weight is not None:
        if packed_weight.numel() != lib.dcnv2_packed_weight_bytes(Cin, Cout) or packed_weight.device != input.device:
            raise RuntimeError("packed_weight does not belong to a [%d,%d,3,3] weight on %s" % (Cout, Cin, input.device))
        flags |= FLAG_PACKED_WEIGHT
    with torch.cuda.device(input.device):
        rc = lib.dcnv2_forward(_lib.ptr(input), _lib.ptr(weight if packed_weight is None else packed_weight), _lib.ptr(bias), _lib.ptr(offset), _lib.ptr(mask),
                               _lib.ptr(out), B, Cin, H, W, Cout, kh, kw, s[0], s[1], p[0], p[1], d[0], d[1],
                               groups, deformable_groups, _lib.ptr(scale), _lib.ptr(shift), flags,
                               _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(input.device))
    _lib.check(rc, "dcnv2_forward")
    return out


def dcnv2_forward_raw27(input, offset_mask, weight, bias=None, offset_mask_bias=None, stride=1, padding=1, dilation=1,
                        scale=None, shift=None, relu=False, fp32_blend=False, packed_weight=None):
    """dcnv2_forward fed by the RAW 27-channel output of `conv_offset_mask` (C ABI dcnv2_forward_raw27): the kernel
    itself takes channels 0..17 as the offsets, applies the sigmoid to channels 18..26 (resnet.py:308-311) and adds
    `offset_mask_bias` when given, so the chunk / cat / sigmoid / bias-add passes over HBM disappear."""
    if not input.is_cuda:
        raise NotImplementedError("Deformable Conv is not supported on CPUs!")    # deform_conv.py:203-204
    lib = _lib.load()
    input = _lib.f32c(input, "input")
    weight = _lib.f32c(weight, "weight")
    offset_mask = _lib.f32c(offset_mask, "offset_mask")
    bias = None if bias is None else _lib.f32c(bias, "bias")
    om_bias = None if offset_mask_bias is None else _lib.f32c(offset_mask_bias, "offset_mask_bias")
    scale = None if scale is None else _lib.f32c(scale, "scale")
    shift = None if shift is None else _lib.f32c(shift, "shift")
    B, Cin, H, W = input.shape
    Cout, cin_w, kh, kw = weight.shape
    if (kh, kw) != (3, 3) or cin_w != Cin:
        raise NotImplementedError("dcnv2_forward_raw27: 3x3, one group")
    Ho = (H + 2 * padding - (dilation * 2 + 1)) // stride + 1
    Wo = (W + 2 * padding - (dilation * 2 + 1)) // stride + 1
    if tuple(offset_mask.shape) != (B, 27, Ho, Wo):
        raise RuntimeError("offset_mask must be [%d,27,%d,%d], got %s" % (B, Ho, Wo, tuple(offset_mask.shape)))
    if om_bias is not None and om_bias.numel() != 27:
        raise RuntimeError("offset_mask_bias must have 27 elements")
    out = torch.empty((B, Cout, Ho, Wo), dtype=torch.float32, device=input.device)
    if out.numel() == 0:
        return out
    nbytes = lib.dcnv2_workspace_bytes(B, Cin, H, W, Cout)
    ws = _workspace(max(nbytes, 16), input.device)
    flags = (FLAG_RELU if relu else 0) | (FLAG_FP32_BLEND if fp32_blend else 0)
    if packed_weight is not None:
        if packed_weight.numel() != lib.dcnv2_packed_weight_bytes(Cin, Cout) or packed_weight.device != input.device:
            raise RuntimeError("packed_weight does not belong to a [%d,%d,3,3] weight on %s" % (Cout, Cin, input.device))
        flags |= FLAG_PACKED_WEIGHT
    with torch.cuda.device(input.device):
        rc = lib.dcnv2_forward_raw27(_lib.ptr(input), _lib.ptr(weight if packed_weight is None else packed_weight), _lib.ptr(bias), _lib.ptr(offset_mask),
                                     _lib.ptr(om_bias), _lib.ptr(out), B, Cin, H, W, Cout, stride, padding, dilation,
                                     _lib.ptr(scale), _lib.ptr(shift), flags,
                                     _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(input.device))
    _lib.check(rc, "dcnv2_forward_raw27")
    return out


def dcnv2_forward_fused_offsets(input, weight, bias, om_weight, om_bias=None, scale=None, shift=None, relu=False,
                                fp32_blend=False, packed_weight=None, return_offset_mask=False):
    """`DCN.forward` of deform_conv.py:505-513 in one call (C ABI dcnv2_forward_fused_offsets): `conv_offset_mask` (`om_weight`
    [27,Cin,3,3], `om_bias` [27]) runs on the tensor cores over the same NHWC bf16 copy of `input` the deformable
    convolution samples, and its raw output goes straight into the DCN kernel (chunk / cat / sigmoid / bias inside).
    stride = padding = dilation = 1.  return_offset_mask: also return the raw 27-channel output (without bias)."""
    if not input.is_cuda:
        raise NotImplementedError("Deformable Conv is not supported on CPUs!")    # deform_conv.py:203-204
    lib = _lib.load()
    input = _lib.f32c(input, "input")
    weight = _lib.f32c(weight, "weight")
    om_weight = _lib.f32c(om_weight, "om_weight")
    bias = None if bias is None else _lib.f32c(bias, "bias")
    om_bias = None if om_bias is None else _lib.f32c(om_bias, "om_bias")
    scale = None if scale is None else _lib.f32c(scale, "scale")
    shift = None if shift is None else _lib.f32c(shift, "shift")
    B, Cin, H, W = input.shape
    Cout, cin_w, kh, kw = weight.shape
    if (kh, kw) != (3, 3) or cin_w != Cin or tuple(om_weight.shape) != (27, Cin, 3, 3):
        raise NotImplementedError("dcnv2_forward_fused_offsets: 3x3, one group, conv_offset_mask [27,Cin,3,3]")
    if om_bias is not None and om_bias.numel() != 27:
        raise RuntimeError("om_bias must have 27 elements")
    out = torch.empty((B, Cout, H, W), dtype=torch.float32, device=input.device)
    raw = torch.empty((B, 27, H, W), dtype=torch.float32, device=input.device) if return_offset_mask else None
    if out.numel() == 0:
        return (out, raw) if return_offset_mask else out
    nbytes = lib.dcnv2_fused_offsets_workspace_bytes(B, Cin, H, W, Cout)
    ws = _workspace(max(nbytes, 16), input.device)
    flags = (FLAG_RELU if relu else 0) | (FLAG_FP32_BLEND if fp32_blend else 0)
    if packed_weight is not None:
        if packed_weight.numel() != lib.dcnv2_packed_weight_bytes(Cin, Cout) or packed_weight.device != input.device:
            raise RuntimeError("packed_weight does not belong to a [%d,%d,3,3] weight on %s" % (Cout, Cin, input.device))
        flags |= FLAG_PACKED_WEIGHT
    with torch.cuda.device(input.device):
        rc = lib.dcnv2_forward_fused_offsets(_lib.ptr(input), _lib.ptr(weight if packed_weight is None else packed_weight),
                                             _lib.ptr(bias), _lib.ptr(om_weight), _lib.ptr(om_bias), _lib.ptr(out), _lib.ptr(raw),
                                             B, Cin, H, W, Cout, _lib.ptr(scale), _lib.ptr(shift), flags,
                                             _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(input.device))
    _lib.check(rc, "dcnv2_forward_fused_offsets")
    return (out, raw) if return_offset_mask else out


def conv3x3(input, weight, bias=None, relu=False):
    """`F.conv2d(input, weight, bias, stride=1, padding=1)` (+ ReLU) for a 3x3 kernel on the tensor cores (C ABI
    ctdet_conv3x3: the DCN forward kernel in its plain mode; bf16 operands, fp32 accumulation) -- the first layer of the ctdet
    heads, centernet.py:112-121.  Inference only (no autograd)."""
    if not input.is_cuda:
        raise NotImplementedError("conv3x3 runs on CUDA tensors only")
    lib = _lib.load()
    input = _lib.f32c(input, "input")
    weight = _lib.f32c(weight, "weight")
    bias = None if bias is None else _lib.f32c(bias, "bias")
    B, Cin, H, W = input.shape
    Cout = weight.shape[0]
    if tuple(weight.shape) != (Cout, Cin, 3, 3) or Cin % 64 != 0:
        raise NotImplementedError("conv3x3: weight [Cout,Cin,3,3] with Cin % 64 == 0")
    out = torch.empty((B, Cout, H, W), dtype=torch.float32, device=input.device)
    if out.numel() == 0:
        return out
    nbytes = lib.ctdet_conv3x3_workspace_bytes(B, Cin, H, W, Cout)
    ws = _workspace(max(nbytes, 16), input.device)
    with torch.cuda.device(input.device):
        rc = lib.ctdet_conv3x3(_lib.ptr(input), _lib.ptr(weight), _lib.ptr(bias), _lib.ptr(out), B, Cin, H, W, Cout,
                               FLAG_RELU if relu else 0, _lib.ptr(ws), ctypes.c_size_t(ws.numel()),
                               _lib.stream_ptr(input.device))
    _lib.check(rc, "ctdet_conv3x3")
    return out


def conv3x3_multi(input, weights, biases, relu=False):
    """Several `conv3x3` of ONE input (C ABI ctdet_conv3x3_multi: the input is packed to NHWC bf16 once) -- the first layers
    of the ctdet heads' branches.  weights: list of [Cout_i,Cin,3,3]; biases: list of [Cout_i] or None.  Returns a list."""
    if not input.is_cuda:
        raise NotImplementedError("conv3x3 runs on CUDA tensors only")
    lib = _lib.load()
    input = _lib.f32c(input, "input")
    weights = [_lib.f32c(w, "weight") for w in weights]
    biases = [None if b is None else _lib.f32c(b, "bias") for b in biases]
    B, Cin, H, W = input.shape
    n = len(weights)
    for w in weights:
        if w.dim() != 4 or tuple(w.shape[1:]) != (Cin, 3, 3) or Cin % 64 != 0:
            raise NotImplementedError("conv3x3: weight [Cout,Cin,3,3] with Cin % 64 == 0")
    outs = [torch.empty((B, w.shape[0], H, W), dtype=torch.float32, device=input.device) for w in weights]
    if B == 0 or H * W == 0:
        return outs
    cmax = max(w.shape[0] for w in weights)
    nbytes = lib.ctdet_conv3x3_workspace_bytes(B, Cin, H, W, cmax)
    ws = _workspace(max(nbytes, 16), input.device)
    vp = ctypes.c_void_p * n
    wp = vp(*[w.data_ptr() for w in weights])
    bp = vp(*[(b.data_ptr() if b is not None else None) for b in biases])
    op = vp(*[o.data_ptr() for o in outs])
    cp = (ctypes.c_int * n)(*[w.shape[0] for w in weights])
    with torch.cuda.device(input.device):
        rc = lib.ctdet_conv3x3_multi(_lib.ptr(input), n, ctypes.cast(wp, ctypes.c_void_p), ctypes.cast(bp, ctypes.c_void_p),
                                     ctypes.cast(op, ctypes.c_void_p), ctypes.cast(cp, ctypes.c_void_p), B, Cin, H, W,
                                     FLAG_RELU if relu else 0, _lib.ptr(ws), ctypes.c_size_t(ws.numel()),
                                     _lib.stream_ptr(input.device))
    _lib.check(rc, "ctdet_conv3x3_multi")
    return outs


def dcnv2_backward(input, offset, mask, weight, grad_output, stride=1, padding=0, dilation=1, groups=1,
                   deformable_groups=1, with_bias=True, need_input_grads=True, need_weight_grad=True):
    """Low-level call: returns (grad_input, grad_offset, grad_mask, grad_weight, grad_bias); entries that were
    not requested are None.  Mirrors `_ModulatedDeformConv.backward` (deform_conv.py:225-262)."""
    if not grad_output.is_cuda:
        raise NotImplementedError("Deformable Conv is not supported on CPUs!")
    lib = _lib.load()
    input, weight = _lib.f32c(input, "input"), _lib.f32c(weight, "weight")
    offset, mask = _lib.f32c(offset, "offset"), _lib.f32c(mask, "mask")
    grad_output = _lib.f32c(grad_output, "grad_output")
    s, p, d = _pair(stride), _pair(padding), _pair(dilation)
    B, Cin, H, W = input.shape
    Cout, _, kh, kw = weight.shape
    Ho, Wo = grad_output.shape[2], grad_output.shape[3]
    if tuple(grad_output.shape) != (B, Cout, offset.shape[2], offset.shape[3]):
        raise RuntimeError("grad_output shape %s does not match the forward output" % (tuple(grad_output.shape),))
    dev = input.device
    gi = go = gm = gw = gb = None
    if need_input_grads:
        gi, go, gm = torch.empty_like(input), torch.empty_like(offset), torch.empty_like(mask)
    if need_weight_grad:
        gw = torch.empty_like(weight)
        gb = torch.empty(Cout, dtype=torch.float32, device=dev) if with_bias else None
    if B == 0 or grad_output.numel() == 0:
        return tuple(None if t is None else t.zero_() for t in (gi, go, gm, gw, gb))
    nbytes = lib.dcnv2_backward_workspace_bytes(B, Cin, H, W, Cout, Ho, Wo)
    ws = _workspace(max(nbytes, 16), dev)
    with torch.cuda.device(dev):
        rc = lib.dcnv2_backward(_lib.ptr(input), _lib.ptr(weight), _lib.ptr(offset), _lib.ptr(mask), _lib.ptr(grad_output),
                                _lib.ptr(gi), _lib.ptr(gw), _lib.ptr(gb), _lib.ptr(go), _lib.ptr(gm),
                                B, Cin, H, W, Cout, kh, kw, s[0], s[1], p[0], p[1], d[0], d[1], groups, deformable_groups, 0,
                                _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr(dev))
    _lib.check(rc, "dcnv2_backward")
    return gi, go, gm, gw, gb


class _ModulatedDeformConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, offset, mask, weight, bias=None, stride=1, padding=0, dilation=1, groups=1,
                deformable_groups=1):
        ctx.cfg = (stride, padding, dilation, groups, deformable_groups)
        ctx.with_bias = bias is not None
        if weight.requires_grad or mask.requires_grad or offset.requires_grad or input.requires_grad:
            if weight.shape[0] % 64 != 0:
                # the backward kernels tile Cout by 64 (the forward takes multiples of 16): fail here, not inside
                # autograd's backward
                raise NotImplementedError("modulated_deform_conv: training needs out_channels %% 64 == 0 (got %d); "
                                          "the forward alone supports multiples of 16" % weight.shape[0])
            ctx.save_for_backward(input, offset, mask, weight)               # deform_conv.py:211
        return dcnv2_forward(input, offset, mask, weight, bias, stride, padding, dilation, groups, deformable_groups)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_output):
        input, offset, mask, weight = ctx.saved_tensors
        stride, padding, dilation, groups, dg = ctx.cfg
        need_in = any(ctx.needs_input_grad[:3])
        need_w = ctx.needs_input_grad[3] or (ctx.with_bias and ctx.needs_input_grad[4])
        gi, go, gm, gw, gb = dcnv2_backward(input, offset, mask, weight, grad_output, stride, padding, dilation, groups,
                                            dg, ctx.with_bias, need_in, need_w)
        return gi, go, gm, gw, gb, None, None, None, None, None


modulated_deform_conv = _ModulatedDeformConv.apply


class ModulatedDeformConv(nn.Module):
    """deform_conv.py:406-495."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 deformable_groups=1, bias=True, norm=None, activation=None):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size = _pair(kernel_size)
        self.stride, self.padding, self.dilation = stride, padding, dilation
        self.groups, self.deformable_groups = groups, deformable_groups
        self.with_bias = bias
        self.norm, self.activation = norm, activation
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels // groups, *self.kernel_size))
        self.bias = nn.Parameter(torch.empty(out_channels)) if bias else None
        nn.init.kaiming_uniform_(self.weight, nonlinearity="relu")
        if self.bias is not None:
            nn.init.constant_(self.bias, 0)

    def forward(self, x, offset, mask):
        x = modulated_deform_conv(x, offset, mask, self.weight, self.bias, self.stride, self.padding,
                                  self.dilation, self.groups, self.deformable_groups)
        if self.norm is not None:
            x = self.norm(x)
        if self.activation is not None:
            x = self.activation(x)
        return x


class DCN(ModulatedDeformConv):
    """The `dcn_v2.DCN` module DeformConvV2 expects (submodule absent from the reference tree):
    a zero-initialised 27-channel conv predicts (dh, dw) offsets and the sigmoid mask."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation=1, deformable_groups=1):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, 1, deformable_groups)
        k = self.kernel_size
        self.conv_offset_mask = nn.Conv2d(in_channels, deformable_groups * 3 * k[0] * k[1], kernel_size=k,
                                          stride=_pair(stride), padding=_pair(padding), bias=True)
        n = in_channels * k[0] * k[1]
        stdv = 1.0 / math.sqrt(n)
        self.weight.data.uniform_(-stdv, stdv)
        self.bias.data.zero_()
        self.conv_offset_mask.weight.data.zero_()
        self.conv_offset_mask.bias.data.zero_()

    def offset_mask(self, x):
        o1, o2, mask = torch.chunk(self.conv_offset_mask(x), 3, dim=1)      # resnet.py:308-311
        return torch.cat((o1, o2), dim=1), torch.sigmoid(mask)

    def forward(self, x):
        offset, mask = self.offset_mask(x)
        return modulated_deform_conv(x, offset, mask, self.weight, self.bias, self.stride, self.padding,
                                     self.dilation, self.groups, self.deformable_groups)


class DeformConvV2(nn.Module):
    """deform_conv.py:498-519: DCN followed by BatchNorm + ReLU.  In eval mode the BatchNorm is
    folded into the kernel's epilogue (one pass instead of three)."""

    def __init__(self, chi, cho):
        super().__init__()
        self.actf = nn.Sequential(nn.BatchNorm2d(cho, momentum=0.1), nn.ReLU(inplace=True))
        self.conv = DCN(chi, cho, kernel_size=(3, 3), stride=1, padding=1, dilation=1, deformable_groups=1)
        nn.init.uniform_(self.actf[0].weight.data)
        self.fuse_offset_conv = True      # eval: conv_offset_mask on the tensor cores inside the DCN call (bf16 operands)

    def _packed_weight(self):
        """Inference: the tensor-core image of the DCN weights, rebuilt when the parameter is replaced, moved or
        modified in place (`_version`; writes through `.data` do not bump it -- call `.invalidate_packed_weight()`)."""
        w = self.conv.weight
        key = (w.data_ptr(), w._version, str(w.device))
        if getattr(self, "_packed_key", None) != key:
            self._packed = dcnv2_pack_weight(w)
            self._packed_key = key
        return self._packed

    def invalidate_packed_weight(self):
        self._packed_key = None

    def forward(self, x):
        bn = self.actf[0]
        if not self.training and not torch.is_grad_enabled() and bn.track_running_stats:
            scale = bn.weight / torch.sqrt(bn.running_var + bn.eps)
            shift = bn.bias - bn.running_mean * scale
            com = self.conv.conv_offset_mask
            # the 27-channel convolution without its bias; bias, chunk / cat and the mask's sigmoid happen inside
            # the DCN kernel's geometry stage
            if self.fuse_offset_conv and com.stride == (1, 1) and com.padding == (1, 1) and com.dilation == (1, 1):
                # conv_offset_mask on the tensor cores over the DCN's own NHWC bf16 copy of x (one call, x packed once)
                return dcnv2_forward_fused_offsets(x, self.conv.weight, self.conv.bias, com.weight, com.bias, scale=scale,
                                                   shift=shift, relu=True, packed_weight=self._packed_weight())
            raw = torch.nn.functional.conv2d(x, com.weight, None, com.stride, com.padding, com.dilation, com.groups)
            return dcnv2_forward_raw27(x, raw, self.conv.weight, self.conv.bias, com.bias, 1, 1, 1,
                                       scale=scale, shift=shift, relu=True, packed_weight=self._packed_weight())
        return self.actf(self.conv(x))
