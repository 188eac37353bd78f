"""`ctdet_decode_static` -- the allocation-free variant of `ctdet_decode` (include/ctdet_b200.h; no counterpart in the
reference) -- called through the C ABI and compared with the existing
entry point `ctdet_decode` on the same inputs.

Collected LAST on purpose (file name).  Passes alone in a fresh process and after
the other tests (it only shares the library's static scratch array with itself).

Tolerance: none -- the decode is integer / selection work on the same kernels, the outputs are compared bit for bit
(`torch.equal`), as in tests/test_gpu_parity.py.
"""
import ctypes

import pytest
import torch

from detectron2_centernet_b200 import _lib
from detectron2_centernet_b200 import synthetic as S

DEV = "cuda:0"
E_WORKSPACE = 4


def _inputs(seed, B, C, H, W):
    hm, wh, reg = S.make_head_outputs(seed, B, C, H, W, tie_free=True)
    return [torch.from_numpy(a).to(DEV) for a in (hm, wh, reg)]


def _outputs(B, K):
    return (torch.full((B, K, 4), -7.0, dtype=torch.float32, device=DEV),
            torch.full((B, K), -7.0, dtype=torch.float32, device=DEV),
            torch.full((B, K), -7, dtype=torch.int32, device=DEV),
            torch.full((B, K), -7, dtype=torch.int64, device=DEV))


def _decode_existing(lib, hm, wh, reg, K, flags):
    B, C, H, W = hm.shape
    n = lib.ctdet_decode_workspace_bytes(B, C, H, W, K)
    ws = torch.zeros(n, dtype=torch.uint8, device=DEV)
    out = _outputs(B, K)
    rc = lib.ctdet_decode(_lib.ptr(hm), _lib.ptr(wh), _lib.ptr(reg), B, C, H, W, K, 4.0, flags,
                          *[_lib.ptr(o) for o in out], _lib.ptr(ws), ctypes.c_size_t(n), _lib.stream_ptr(DEV))
    assert rc == 0, lib.ctdet_error_string(rc)
    torch.cuda.synchronize()
    return out


def _decode_static(lib, hm, wh, reg, K, flags, out):
    B, C, H, W = hm.shape
    rc = lib.ctdet_decode_static(_lib.ptr(hm), _lib.ptr(wh), _lib.ptr(reg), B, C, H, W, K, 4.0, flags,
                                 *[_lib.ptr(o) for o in out], _lib.stream_ptr(DEV))
    assert rc == 0, (lib.ctdet_error_string(rc), lib.ctdet_last_cuda_error())


@pytest.mark.gpu
def test_static_scratch_decode_equals_ctdet_decode_on_two_calls_in_a_row():
    lib = _lib.load()
    assert lib.ctdet_decode_static_max_bytes() == 16 << 20
    # (shape, K, flags): two calls in a row on the same shape, then other batch sizes and map shapes, so that the
    # key segments of one call sit where the next call's headers / histograms must start at zero; 1 = logits mode
    plan = [((4, 80, 128, 128), 100, 1), ((4, 80, 128, 128), 100, 1), ((2, 80, 128, 128), 100, 0),
            ((7, 6, 32, 32), 20, 1), ((1, 3, 30, 50), 40, 0), ((4, 80, 128, 128), 100, 1)]
    with torch.cuda.device(DEV):
        cases = []
        for i, (shape, K, flags) in enumerate(plan):
            hm, wh, reg = _inputs(500 + i, *shape)
            if not flags:
                hm = torch.clamp(torch.sigmoid(hm), 1e-4, 1 - 1e-4)
            cases.append((hm, wh, reg, K, flags, _decode_existing(lib, hm, wh, reg, K, flags), _outputs(shape[0], K)))
        # first call: also pays CUDA's lazy loading of the kernels; from here on nothing may be allocated
        _decode_static(lib, *cases[0][:5], cases[0][6])
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        torch_bytes0 = torch.cuda.memory_allocated()
        for hm, wh, reg, K, flags, want, got in cases[1:]:          # back to back on one stream, no sync in between
            _decode_static(lib, hm, wh, reg, K, flags, got)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == torch_bytes0
        assert torch.cuda.mem_get_info()[0] == free0, "device memory was allocated by ctdet_decode_static"
    for i, (hm, wh, reg, K, flags, want, got) in enumerate(cases):
        for name, w, g in zip(("boxes", "scores", "classes", "inds"), want, got):
            assert torch.equal(w, g), (i, name)


@pytest.mark.gpu
def test_static_scratch_decode_under_graph_capture_replays():
    """The use case named in the header: captured into a CUDA graph (no allocation is legal there) and replayed."""
    lib = _lib.load()
    with torch.cuda.device(DEV):
        hm, wh, reg = _inputs(611, 3, 80, 128, 128)
        want = _decode_existing(lib, hm, wh, reg, 100, 1)
        got = _outputs(3, 100)
        _decode_static(lib, hm, wh, reg, 100, 1, got)               # set-up (symbol address) outside the capture
        torch.cuda.synchronize()
        for o in got:
            o.fill_(-7)
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                _decode_static(lib, hm, wh, reg, 100, 1, got)
        for _ in range(2):
            g.replay()
        torch.cuda.synchronize()
    for w, o in zip(want, got):
        assert torch.equal(w, o)


def test_static_scratch_decode_rejects_shapes_beyond_the_array():
    """Host-side check, made before any CUDA call: runs without a GPU."""
    lib = _lib.load()
    B, C, H, W, K = 32, 80, 128, 128, 100
    assert lib.ctdet_decode_workspace_bytes(B, C, H, W, K) > lib.ctdet_decode_static_max_bytes()
    rc = lib.ctdet_decode_static(None, None, None, B, C, H, W, K, 4.0, 0, None, None, None, None, None)
    assert rc == E_WORKSPACE
    msg = lib.ctdet_last_cuda_error().decode()
    assert "ctdet_decode_static" in msg and "16777216" in msg and "needs" in msg, msg
    assert lib.ctdet_decode_static(None, None, None, 1, 80, 128, 128, 0, 4.0, 0, None, None, None, None, None) == 1
