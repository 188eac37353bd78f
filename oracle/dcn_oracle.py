"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper of oracle/dcn_oracle.c (see its header)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libdcn_oracle.so")
_lib = None


def _load():
    global _lib
    if _lib is None:
        if not os.path.isfile(_SO):
            subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        _lib = ctypes.CDLL(_SO)
        _lib.dcn_oracle_forward.restype = ctypes.c_int
        _lib.dcn_oracle_forward.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int] * 8 + [ctypes.c_void_p]
        _lib.dcn_oracle_backward.restype = ctypes.c_int
        _lib.dcn_oracle_backward.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int] * 8 + [ctypes.c_void_p] * 5
    return _lib


def dcn_forward(x, offset, mask, weight, bias=None, stride=1, pad=1, dil=1, threads=1):
    """x [B,Cin,H,W], offset [B,18,Ho,Wo], mask [B,9,Ho,Wo], weight [Cout,Cin,3,3] -> [B,Cout,Ho,Wo].
    threads > 1 runs images concurrently (ctypes releases the GIL; the C code is single-threaded
    because this image's gcc has no libgomp)."""
    x, offset, mask, weight = [np.ascontiguousarray(a, np.float32) for a in (x, offset, mask, weight)]
    B, Cin, H, W = x.shape
    Cout = weight.shape[0]
    Ho = (H + 2 * pad - (dil * 2 + 1)) // stride + 1
    Wo = (W + 2 * pad - (dil * 2 + 1)) // stride + 1
    assert offset.shape == (B, 18, Ho, Wo) and mask.shape == (B, 9, Ho, Wo) and weight.shape == (Cout, Cin, 3, 3)
    out = np.empty((B, Cout, Ho, Wo), np.float32)
    bptr = None
    if bias is not None:
        bias = np.ascontiguousarray(bias, np.float32)
        bptr = bias.ctypes.data
    fn = _load().dcn_oracle_forward

    def one(b0, b1):
        return fn(x[b0:b1].ctypes.data, offset[b0:b1].ctypes.data, mask[b0:b1].ctypes.data, weight.ctypes.data,
                  bptr, b1 - b0, Cin, H, W, Cout, stride, pad, dil, out[b0:b1].ctypes.data)
    if threads <= 1 or B == 1:
        rcs = [one(0, B)]
    else:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=threads) as ex:
            rcs = list(ex.map(lambda b: one(b, b + 1), range(B)))
    if any(rcs):
        raise MemoryError("dcn oracle scratch allocation failed")
    return out


def dcn_backward(x, offset, mask, weight, grad_out, stride=1, pad=1, dil=1):
    """-> (grad_input, grad_offset, grad_mask, grad_weight, grad_bias), the five outputs of
    modulated_deform_conv_cuda_backward (deform_conv_cuda.cu:929)."""
    x, offset, mask, weight, grad_out = [np.ascontiguousarray(a, np.float32) for a in (x, offset, mask, weight, grad_out)]
    B, Cin, H, W = x.shape
    Cout = weight.shape[0]
    Ho = (H + 2 * pad - (dil * 2 + 1)) // stride + 1
    Wo = (W + 2 * pad - (dil * 2 + 1)) // stride + 1
    assert offset.shape == (B, 18, Ho, Wo) and mask.shape == (B, 9, Ho, Wo) and grad_out.shape == (B, Cout, Ho, Wo)
    gx, go, gm = np.empty_like(x), np.empty_like(offset), np.empty_like(mask)
    gw, gb = np.empty_like(weight), np.empty((Cout,), np.float32)
    rc = _load().dcn_oracle_backward(x.ctypes.data, offset.ctypes.data, mask.ctypes.data, weight.ctypes.data,
                                     grad_out.ctypes.data, B, Cin, H, W, Cout, stride, pad, dil,
                                     gx.ctypes.data, go.ctypes.data, gm.ctypes.data, gw.ctypes.data, gb.ctypes.data)
    if rc:
        raise MemoryError("dcn oracle scratch allocation failed")
    return gx, go, gm, gw, gb
