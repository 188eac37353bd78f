"""The callers around the DCNv2 contraction: the DLA-34 up-sampling neck and the ctdet heads, as a
host-side mirror of detectron2/modeling/backbone/dla.py (`fill_up_weights` :33, `IDAUp` :152,
`DLAUp` :180, the neck half of `DLA34` :287-315) and of the head construction in
detectron2/modeling/meta_arch/centernet.py:112-134.  Sub-module and parameter names are the reference's
(`proj_i` / `up_i` / `node_i`, `ida_i`, `dla_up`, `ida_up`, `hm` / `wh` / `reg`), so a reference
state_dict loads unchanged; every `DeformConvV2` is the B200 one (tcgen05 kernels, BN + ReLU folded
into the epilogue in eval mode).  The DLA base network (plain cuDNN convolutions) is not part of the
hot path and stays whatever torch module the caller has: this neck takes its six feature maps.
"""
import math

import torch
from torch import nn

from . import _lib
from .deform_conv import DeformConvV2


def _is_ida_up(up):
    """True for the module IDAUp builds at dla.py:166-168: depth-wise ConvTranspose2d(o, o, 2f, stride f,
    padding f // 2), no bias, f a power of two."""
    if not isinstance(up, nn.ConvTranspose2d) or up.bias is not None:
        return False
    f = up.stride[0]
    return (up.stride == (f, f) and up.kernel_size == (2 * f, 2 * f) and up.padding == (f // 2, f // 2) and
            up.output_padding == (0, 0) and up.dilation == (1, 1) and up.groups == up.in_channels == up.out_channels and
            2 <= f <= 16 and (f & (f - 1)) == 0)


def _upsample_add_fwd(x, weight, skip, f):
    lib = _lib.load()
    x = _lib.f32c(x, "x")
    B, C, h, w = x.shape
    if skip is not None:
        skip = _lib.f32c(skip, "skip")
        if tuple(skip.shape) != (B, C, h * f, w * f):
            raise RuntimeError("skip must be [%d,%d,%d,%d], got %s" % (B, C, h * f, w * f, tuple(skip.shape)))
    out = torch.empty((B, C, h * f, w * f), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        rc = lib.ctdet_upsample_add(_lib.ptr(x), _lib.ptr(_lib.f32c(weight, "weight")), _lib.ptr(skip),
                                    B, C, h, w, f, _lib.ptr(out), _lib.stream_ptr(x.device))
    _lib.check(rc, "ctdet_upsample_add")
    return out


class _UpsampleAdd(torch.autograd.Function):
    """`up(x) + skip` with its backward on the GPU kernels too (C ABI ctdet_upsample_add_backward): grad_x and
    grad_weight in one pass over grad_out, grad_skip = grad_out.  In the reference this is the grouped
    ConvTranspose2d's cuDNN forward / backward-data / backward-filter plus the add (dla.py:175-177)."""

    @staticmethod
    def forward(ctx, x, weight, skip, f):
        ctx.f = f
        ctx.has_skip = skip is not None
        ctx.save_for_backward(x, weight)
        return _upsample_add_fwd(x, weight.detach(), skip, f)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_out):
        x, weight = ctx.saved_tensors
        lib = _lib.load()
        f = ctx.f
        grad_out = _lib.f32c(grad_out, "grad_out")
        B, C, h, w = x.shape
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        gw = torch.empty_like(weight) if ctx.needs_input_grad[1] else None
        if gx is not None or gw is not None:
            with torch.cuda.device(x.device):
                rc = lib.ctdet_upsample_add_backward(_lib.ptr(grad_out), _lib.ptr(_lib.f32c(x, "x")),
                                                     _lib.ptr(_lib.f32c(weight, "weight")), B, C, h, w, f, _lib.ptr(gx),
                                                     _lib.ptr(gw), _lib.stream_ptr(x.device))
            _lib.check(rc, "ctdet_upsample_add_backward")
        return gx, gw, (grad_out if ctx.has_skip and ctx.needs_input_grad[2] else None), None


def upsample_add(up, x, skip=None):
    """`up(x) + skip` of IDAUp.forward (dla.py:175-177) in one pass over HBM (C ABI ctdet_upsample_add); with autograd
    the backward runs on ctdet_upsample_add_backward (f = 2 / 4, what DLA-34 uses).  For a module that is not
    IDAUp's up-sampling (or f = 8 / 16 under autograd) the reference's own two torch ops run."""
    fused = x.is_cuda and x.dtype == torch.float32 and _is_ida_up(up) and up.weight.dtype == torch.float32
    needs_grad = torch.is_grad_enabled() and (x.requires_grad or up.weight.requires_grad or
                                              (skip is not None and skip.requires_grad))
    if fused and needs_grad and up.stride[0] not in (2, 4):
        fused = False
    if not fused:
        y = up(x)
        return y if skip is None else y + skip
    if needs_grad:
        return _UpsampleAdd.apply(x, up.weight, skip, up.stride[0])
    return _upsample_add_fwd(x, up.weight.detach(), skip, up.stride[0])


def fill_up_weights(up):
    """dla.py:33-42: the fixed bilinear kernel of the depth-wise transposed convolution."""
    w = up.weight.data
    k = w.size(2)
    f = math.ceil(k / 2)
    c = (2 * f - 1 - f % 2) / (2.0 * f)
    for i in range(k):
        for j in range(w.size(3)):
            w[0, 0, i, j] = (1 - math.fabs(i / f - c)) * (1 - math.fabs(j / f - c))
    w[1:, 0] = w[0, 0]


class IDAUp(nn.Module):
    """dla.py:152-177.  For every level after the first: project to `o` channels (DCN), up-sample by
    `up_f[i]` (depth-wise ConvTranspose2d with the bilinear kernel), add the previous level, and mix
    (`node`, another DCN).  `dcn` lets tests swap in another DeformConvV2 implementation."""

    def __init__(self, o, channels, up_f, dcn=DeformConvV2):
        super().__init__()
        for i in range(1, len(channels)):
            f = int(up_f[i])
            setattr(self, "proj_%d" % i, dcn(channels[i], o))
            up = nn.ConvTranspose2d(o, o, f * 2, stride=f, padding=f // 2, output_padding=0, groups=o, bias=False)
            fill_up_weights(up)
            setattr(self, "up_%d" % i, up)
            setattr(self, "node_%d" % i, dcn(o, o))

    def forward(self, layers, startp, endp):
        for i in range(startp + 1, endp):
            k = i - startp
            x = upsample_add(getattr(self, "up_%d" % k), getattr(self, "proj_%d" % k)(layers[i]), layers[i - 1])
            layers[i] = getattr(self, "node_%d" % k)(x)


class DLAUp(nn.Module):
    """dla.py:180-203: one IDAUp per resolution, deepest first; returns the fused maps, finest first."""

    def __init__(self, startp, channels, scales, in_channels=None, dcn=DeformConvV2):
        super().__init__()
        self.startp = startp
        channels = list(channels)
        in_channels = list(channels if in_channels is None else in_channels)
        self.channels = channels
        scales = [int(s) for s in scales]
        for i in range(len(channels) - 1):
            j = -i - 2
            setattr(self, "ida_%d" % i, IDAUp(channels[j], in_channels[j:], [s // scales[j] for s in scales[j:]], dcn=dcn))
            scales[j + 1:] = [scales[j]] * len(scales[j + 1:])
            in_channels[j + 1:] = [channels[j]] * len(in_channels[j + 1:])

    def forward(self, layers):
        layers = list(layers)
        out = [layers[-1]]
        for i in range(len(layers) - self.startp - 1):
            getattr(self, "ida_%d" % i)(layers, len(layers) - i - 2, len(layers))
            out.insert(0, layers[-1])
        return out


class DLA34Neck(nn.Module):
    """The `dla_up` + `ida_up` half of DLA34 (dla.py:287-315) over the base network's feature maps
    (channels 16, 32, 64, 128, 256, 512 at strides 1 .. 32): 16 DeformConvV2 layers, output = the
    stride-`down_ratio` map with `channels[first_level]` channels (last element of the returned list)."""

    def __init__(self, channels=(16, 32, 64, 128, 256, 512), down_ratio=4, last_level=5, dcn=DeformConvV2):
        super().__init__()
        assert down_ratio in (2, 4, 8, 16)
        self.channels = list(channels)
        self.first_level = int(math.log2(down_ratio))
        self.last_level = last_level
        up_channels = self.channels[self.first_level:]
        self.dla_up = DLAUp(self.first_level, up_channels, [2 ** i for i in range(len(up_channels))], dcn=dcn)
        self.ida_up = IDAUp(self.channels[self.first_level], self.channels[self.first_level:self.last_level],
                            [2 ** i for i in range(self.last_level - self.first_level)], dcn=dcn)

    def forward(self, base_features):
        x = self.dla_up(base_features)
        y = [x[i].clone() for i in range(self.last_level - self.first_level)]
        self.ida_up(y, 0, len(y))
        return y


class CtdetHead(nn.Module):
    """centernet.py:112-134: per head `Conv3x3(in, head_conv) + ReLU + Conv(final_kernel)(head_conv, n)`;
    the `hm` bias starts at -2.19 (:122-123), the other heads' biases at 0 (`fill_fc_weights`).
    Returns {'hm', 'wh', 'reg'} (names pinned by export/meta_modeling.py:200)."""

    def __init__(self, in_channels=64, num_classes=80, head_conv=256, final_kernel=1, heads=None):
        super().__init__()
        self.heads = dict(heads) if heads is not None else {"hm": num_classes, "wh": 2, "reg": 2}
        for name, n in self.heads.items():
            if head_conv > 0:
                fc = nn.Sequential(nn.Conv2d(in_channels, head_conv, kernel_size=3, padding=1, bias=True),
                                   nn.ReLU(inplace=True),
                                   nn.Conv2d(head_conv, n, kernel_size=final_kernel, stride=1,
                                             padding=final_kernel // 2, bias=True))
                last = fc[-1]
            else:
                fc = nn.Conv2d(in_channels, n, kernel_size=final_kernel, stride=1, padding=final_kernel // 2, bias=True)
                last = fc
            if "hm" in name.lower():
                last.bias.data.fill_(-2.19)
            else:
                for m in fc.modules():
                    if isinstance(m, nn.Conv2d) and m.bias is not None:
                        nn.init.constant_(m.bias, 0)
            setattr(self, name.lower(), fc)

    fuse_first_conv = True     # eval, no autograd: the 3x3 convolution + ReLU of every branch on the tensor cores (bf16 operands)

    def _fusable_trunk(self, fc, x):
        c = fc[0] if isinstance(fc, nn.Sequential) else None
        return (self.fuse_first_conv and not torch.is_grad_enabled() and x.is_cuda and isinstance(c, nn.Conv2d) and
                c.kernel_size == (3, 3) and c.stride == (1, 1) and c.padding == (1, 1) and c.dilation == (1, 1) and
                c.groups == 1 and c.in_channels % 64 == 0 and len(fc) == 3 and isinstance(fc[1], nn.ReLU))

    def _trunks(self, x):
        """Every branch but its last layer -- Conv3x3 + ReLU (centernet.py:112-121) -- for all branches: one
        `conv3x3_multi` call (x packed once) where the branches allow it, the torch modules otherwise."""
        names = [n.lower() for n in self.heads]
        fcs = {n: getattr(self, n) for n in names}
        fused = [n for n in names if self._fusable_trunk(fcs[n], x)]
        out = {}
        if fused:
            from .deform_conv import conv3x3_multi
            feats = conv3x3_multi(x, [fcs[n][0].weight for n in fused], [fcs[n][0].bias for n in fused], relu=True)
            out.update(zip(fused, feats))
        for n in names:
            if n not in out:
                out[n] = fcs[n][:-1](x) if isinstance(fcs[n], nn.Sequential) else x
        return out

    def _last(self, name, feat):
        fc = getattr(self, name.lower())
        return fc[-1](feat) if isinstance(fc, nn.Sequential) else fc(feat)

    def forward(self, x):
        feats = self._trunks(x)
        return {name: self._last(name, feats[name.lower()]) for name in self.heads}

    def decode(self, x, down_ratio=4, K=100, fuse=True):
        """Eval path `ctdet_decode(hm.sigmoid_(), wh, reg, K)` (centernet.py:151-166) on the heads of `x` for a batch: with
        `fuse` and shapes `decode.head_decode_fusable` takes, the hm branch's final 1x1 convolution runs inside the decode
        (`hm_head_decode_batch`: the logits never reach HBM); otherwise forward() + ctdet_decode_batch(logits=True).
        Returns boxes, scores, classes, inds."""
        from . import decode as D
        hm = self.hm
        feats = self._trunks(x)
        wh = self._last("wh", feats["wh"])
        reg = self._last("reg", feats["reg"]) if "reg" in feats else None
        last = hm[-1] if isinstance(hm, nn.Sequential) else None
        if (fuse and last is not None and last.kernel_size == (1, 1) and
                D.head_decode_fusable(last.in_channels, last.out_channels, x.shape[2], x.shape[3])):
            return D.hm_head_decode_batch(feats["hm"], last.weight, last.bias, wh, reg, down_ratio=down_ratio, K=K)
        return D.ctdet_decode_batch(self._last("hm", feats["hm"]), wh, reg, down_ratio=down_ratio, K=K, logits=True)
