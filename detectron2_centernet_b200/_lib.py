"""ctypes binding of libctdet_b200.so (the C ABI declared in include/ctdet_b200.h).

There is NO fallback: if the shared library is missing or a tensor is not a CUDA tensor the call
raises.  Error convention mirrors the reference's pybind ops (detectron2/layers/deform_conv.py:
203-204 raises NotImplementedError for CPU tensors; TORCH_CHECK failures surface as RuntimeError,
deform_conv_cuda.cu:824-848).
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libctdet_b200.so")
CSRC = os.path.join(_HERE, "csrc")

_lib = None

c_int, c_float, c_size_t, c_void_p = ctypes.c_int, ctypes.c_float, ctypes.c_size_t, ctypes.c_void_p

# name -> (restype, argtypes); must list every symbol include/ctdet_b200.h declares
SIGNATURES = {
    "ctdet_abi_version": (c_int, []),
    "ctdet_error_string": (ctypes.c_char_p, [c_int]),
    "ctdet_last_cuda_error": (ctypes.c_char_p, []),
    "ctdet_target_gen": (c_int, [c_void_p] * 3 + [c_int] * 4 + [c_void_p] * 6 + [c_void_p]),
    "ctdet_loss_workspace_bytes": (c_size_t, []),
    "ctdet_loss_fwd_bwd": (c_int, [c_void_p] * 11 + [c_int] * 6 + [c_float] * 3 + [c_void_p] * 4 +
                           [c_void_p, c_size_t, c_void_p]),
    "ctdet_fused_workspace_bytes": (c_size_t, [c_int] * 2),
    "ctdet_target_loss_fused": (c_int, [c_void_p] * 8 + [c_int] * 5 + [c_float] * 3 + [c_void_p] * 9 +
                                [c_void_p, c_size_t, c_void_p]),
    "ctdet_head_focal_fused": (c_int, [c_void_p] * 7 + [c_int] * 6 + [c_float] + [c_void_p] * 3 + [c_size_t, c_void_p]),
    "ctdet_head_focal_fused_bwd": (c_int, [c_void_p] * 7 + [c_int] * 6 + [c_float] + [c_void_p] * 6 + [c_size_t, c_void_p]),
    "ctdet_scale_grad": (c_int, [c_void_p, c_size_t, c_void_p, c_void_p]),
    "ctdet_decode_workspace_bytes": (c_size_t, [c_int] * 5),
    "ctdet_decode": (c_int, [c_void_p] * 3 + [c_int] * 5 + [c_float, c_int] + [c_void_p] * 4 +
                     [c_void_p, c_size_t, c_void_p]),
    "ctdet_decode_static_max_bytes": (c_size_t, []),
    "ctdet_decode_static": (c_int, [c_void_p] * 3 + [c_int] * 5 + [c_float, c_int] + [c_void_p] * 4 + [c_void_p]),
    "ctdet_head_decode_fused_workspace_bytes": (c_size_t, [c_int] * 6),
    "ctdet_head_decode_fused": (c_int, [c_void_p] * 5 + [c_int] * 6 + [c_float, c_int] + [c_void_p] * 5 +
                                [c_void_p, c_size_t, c_void_p]),
    "ctdet_nms": (c_int, [c_void_p] + [c_int] * 4 + [c_void_p, c_void_p]),
    "ctdet_gather_feat": (c_int, [c_void_p] * 2 + [c_int] * 4 + [c_void_p, c_void_p]),
    "ctdet_postprocess": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_float, c_void_p, c_void_p]),
    "ctdet_preprocess": (c_int, [c_void_p, c_void_p] + [c_int] * 5 + [c_void_p] * 4),
    "ctdet_upsample_add": (c_int, [c_void_p] * 3 + [c_int] * 5 + [c_void_p, c_void_p]),
    "ctdet_upsample_add_backward": (c_int, [c_void_p] * 3 + [c_int] * 5 + [c_void_p, c_void_p, c_void_p]),
    "dcnv2_workspace_bytes": (c_size_t, [c_int] * 5),
    "dcnv2_packed_weight_bytes": (c_size_t, [c_int] * 2),
    "dcnv2_pack_weight": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "dcnv2_forward": (c_int, [c_void_p] * 6 + [c_int] * 15 + [c_void_p, c_void_p, c_int] +
                      [c_void_p, c_size_t, c_void_p]),
    "dcnv2_fused_offsets_workspace_bytes": (c_size_t, [c_int] * 5),
    "dcnv2_forward_fused_offsets": (c_int, [c_void_p] * 7 + [c_int] * 5 + [c_void_p, c_void_p, c_int] +
                                    [c_void_p, c_size_t, c_void_p]),
    "ctdet_conv3x3_workspace_bytes": (c_size_t, [c_int] * 5),
    "ctdet_conv3x3": (c_int, [c_void_p] * 4 + [c_int] * 6 + [c_void_p, c_size_t, c_void_p]),
    "ctdet_conv3x3_multi": (c_int, [c_void_p, c_int] + [c_void_p] * 4 + [c_int] * 5 + [c_void_p, c_size_t, c_void_p]),
    "dcnv2_forward_raw27": (c_int, [c_void_p] * 6 + [c_int] * 8 + [c_void_p, c_void_p, c_int] +
                            [c_void_p, c_size_t, c_void_p]),
    "dcnv2_backward_workspace_bytes": (c_size_t, [c_int] * 7),
    "dcnv2_backward": (c_int, [c_void_p] * 10 + [c_int] * 16 + [c_void_p, c_size_t, c_void_p]),
}


def build(verbose=False):
    """Compiles the library in-tree with nvcc for sm_100a (csrc/Makefile)."""
    r = subprocess.run(["make", "-C", CSRC], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libctdet_b200.so failed")
    return LIB_PATH


def load():
    """Loads the shared library (once) and declares every prototype."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(
            "libctdet_b200.so not found at %s -- build it with `python __graft_entry__.py` or "
            "`make -C %s`; there is no CPU / PyTorch fallback for the ctdet hot path" % (LIB_PATH, CSRC))
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.ctdet_abi_version() != 1:
        raise RuntimeError("libctdet_b200.so ABI version mismatch")
    _lib = lib
    return lib


_ws_cache = {}


def check(code, what):
    if code == 0:
        return
    # The cached loss / decode workspaces rely on "zero on entry, left zero on exit".  A multi-launch call that failed
    # part-way may have left them dirty: drop them, the next call allocates fresh zeroed ones.
    _ws_cache.clear()
    lib = load()
    msg = lib.ctdet_error_string(code).decode()
    if code == 5:
        msg += ": " + lib.ctdet_last_cuda_error().decode()
    if code == 6:
        raise NotImplementedError("%s: %s" % (what, msg))
    raise RuntimeError("%s: %s" % (what, msg))


def require_cuda(what, *tensors):
    """No CPU path exists (like deform_conv.py:203-204)."""
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NotImplementedError("%s is not supported on CPU tensors (CUDA sm_100a only)" % what)


def ptr(t):
    return None if t is None else c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


def f32c(t, what):
    """fp32 + contiguous, as the reference's kernels require (deform_conv_cuda.cu:824-825)."""
    if t.dtype != torch.float32:
        raise RuntimeError("%s must be float32, got %s" % (what, t.dtype))
    return t if t.is_contiguous() else t.contiguous()


def zero_workspace(key, nbytes, device):
    """A cached, zero-initialised byte buffer (the loss kernel leaves it zeroed after every call)."""
    k = (key, str(device), torch.cuda.current_stream(device).cuda_stream)
    buf = _ws_cache.get(k)
    if buf is None or buf.numel() < nbytes:
        buf = torch.zeros(max(nbytes, 64), dtype=torch.uint8, device=device)
        _ws_cache[k] = buf
    return buf
