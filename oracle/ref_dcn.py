"""TEST INFRASTRUCTURE ONLY — loader and call wrappers for oracle/_ref/ref_dcn*.so, the REFERENCE's own
modulated deformable convolution CUDA code compiled unmodified (oracle/build_ref.py).

The two wrappers restate the reference's Python call site, `_ModulatedDeformConv.forward/backward`
(detectron2/layers/deform_conv.py:180-274): caller-allocated output / zero-initialised gradients, two
empty scratch tensors (`ones`, `columns`) the callee resizes, fake 1-element bias when `bias is None`.

Needs a GPU (the reference has no CPU implementation: deform_conv.h:311).  Imported only by tests/,
__graft_entry__.smoke() and bench.py's `reference_gpu` leg.
"""
import importlib.util
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_mod = None


def available():
    from . import build_ref
    return os.path.isfile(build_ref.out_path())


def load():
    """Loads the prebuilt module (builds it first when the reference tree is present and it is stale)."""
    global _mod
    if _mod is None:
        from . import build_ref
        path = build_ref.build(verbose=False)
        if path is None or not os.path.isfile(path):
            raise RuntimeError("oracle/_ref is not built and /root/reference is absent: run "
                               "`python oracle/build_ref.py` in the container that holds the reference")
        spec = importlib.util.spec_from_file_location(build_ref.MOD, path)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


def _out_hw(H, W, k, stride, padding, dilation):
    return ((H + 2 * padding - (dilation * (k - 1) + 1)) // stride + 1,
            (W + 2 * padding - (dilation * (k - 1) + 1)) // stride + 1)


def forward(input, offset, mask, weight, bias=None, stride=1, padding=0, dilation=1, groups=1,
            deformable_groups=1):
    """deform_conv.py:180-235 -> deform_conv_cuda.cu:804-927.  CUDA fp32 tensors in, output fp32."""
    C = load()
    with_bias = bias is not None
    if not with_bias:
        bias = input.new_empty(1)
    Ho, Wo = _out_hw(input.shape[2], input.shape[3], weight.shape[2], stride, padding, dilation)
    output = input.new_empty((input.shape[0], weight.shape[0], Ho, Wo))
    ones, columns = input.new_empty(0), input.new_empty(0)
    C.modulated_deform_conv_forward(input.contiguous(), weight.contiguous(), bias, ones, offset.contiguous(),
                                    mask.contiguous(), output, columns, weight.shape[2], weight.shape[3],
                                    stride, stride, padding, padding, dilation, dilation, groups,
                                    deformable_groups, with_bias)
    return output


def backward(input, offset, mask, weight, grad_output, bias=None, stride=1, padding=0, dilation=1, groups=1,
             deformable_groups=1):
    """deform_conv.py:237-290 -> deform_conv_cuda.cu:929-1129.
    -> (grad_input, grad_offset, grad_mask, grad_weight, grad_bias or None)."""
    C = load()
    with_bias = bias is not None
    if not with_bias:
        bias = input.new_empty(1)
    grad_input = torch.zeros_like(input)
    grad_offset = torch.zeros_like(offset)
    grad_mask = torch.zeros_like(mask)
    grad_weight = torch.zeros_like(weight)
    grad_bias = torch.zeros_like(bias)
    ones, columns = input.new_empty(0), input.new_empty(0)
    C.modulated_deform_conv_backward(input.contiguous(), weight.contiguous(), bias, ones, offset.contiguous(),
                                     mask.contiguous(), columns, grad_input, grad_weight, grad_bias, grad_offset,
                                     grad_mask, grad_output.contiguous(), weight.shape[2], weight.shape[3],
                                     stride, stride, padding, padding, dilation, dilation, groups,
                                     deformable_groups, with_bias)
    return grad_input, grad_offset, grad_mask, grad_weight, (grad_bias if with_bias else None)
