"""CPU: the C-ABI library builds, loads and exports every symbol include/ctdet_b200.h declares;
host-side argument validation (no kernels are launched without a GPU)."""
import ctypes
import os
import re

import pytest
import torch

from detectron2_centernet_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.isfile(_lib.LIB_PATH):
        _lib.build()
    return _lib.load()


def header_symbols():
    src = open(os.path.join(ROOT, "include", "ctdet_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:ctdet|dcnv2)_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(lib):
    names = header_symbols()
    assert "ctdet_decode" in names and "ctdet_target_gen" in names and "ctdet_loss_fwd_bwd" in names
    for n in names:
        assert hasattr(lib, n), "libctdet_b200.so does not export %s" % n
        assert n in _lib.SIGNATURES, "no ctypes prototype for %s" % n
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_error_strings(lib):
    assert lib.ctdet_abi_version() == 1
    assert lib.ctdet_error_string(0) == b"ok"
    assert b"workspace" in lib.ctdet_error_string(4)
    assert lib.ctdet_loss_workspace_bytes() >= 48


def test_argument_validation_without_gpu(lib):
    null = ctypes.c_void_p(None)
    fake = ctypes.c_void_p(0x1000)
    # decode: K > H*W is what torch.topk rejects at centernet.py:411
    assert lib.ctdet_decode(fake, fake, null, 1, 2, 4, 4, 17, 1.0, 0, fake, fake, fake, null, fake, 1 << 20, null) == 1
    assert lib.ctdet_decode(fake, fake, null, 1, 2, 4, 4, 4, 1.0, 64, fake, fake, fake, null, fake, 1 << 20, null) == 1
    assert lib.ctdet_decode(null, fake, null, 1, 2, 4, 4, 4, 1.0, 0, fake, fake, fake, null, fake, 1 << 20, null) == 2
    assert lib.ctdet_decode(fake, fake, null, 1, 2, 4, 4, 4, 1.0, 0, fake, fake, fake, null, fake, 8, null) == 4
    assert lib.ctdet_decode(ctypes.c_void_p(0x1004), fake, null, 1, 2, 4, 4, 4, 1.0, 0, fake, fake, fake, null, fake, 1 << 20, null) == 3
    assert lib.ctdet_decode(fake, fake, null, 0, 2, 4, 4, 4, 1.0, 0, fake, fake, fake, null, fake, 0, null) == 0   # empty batch
    assert lib.ctdet_decode_workspace_bytes(64, 80, 128, 128, 100) >= 64 * 80 * 100 * 8
    # target gen
    assert lib.ctdet_target_gen(fake, fake, null, 1, 80, 128, 128, fake, fake, fake, fake, fake, null, null) == 2
    assert lib.ctdet_target_gen(fake, fake, fake, 1, 0, 128, 128, fake, fake, fake, fake, fake, null, null) == 1
    assert lib.ctdet_target_gen(fake, fake, fake, 0, 80, 128, 128, fake, fake, fake, fake, fake, null, null) == 0
    # loss: H*W must be a multiple of 4; gradients need aligned pointers
    args = [fake] * 9 + [null, null]
    assert lib.ctdet_loss_fwd_bwd(*args, 1, 2, 3, 3, 128, 0, 1.0, 0.1, 1.0, fake, null, null, null, fake, 64, null) == 6
    assert lib.ctdet_loss_fwd_bwd(*args, 1, 2, 4, 4, 128, 0, 1.0, 0.1, 1.0, fake, null, null, null, fake, 8, null) == 4
    assert lib.ctdet_loss_fwd_bwd(*args, 1, 2, 4, 4, 128, 8, 1.0, 0.1, 1.0, fake, null, null, null, fake, 64, null) == 1


def test_cpu_tensors_are_rejected_not_computed():
    """No CPU fallback: like deform_conv.py:203-204 the ops raise on CPU tensors."""
    from detectron2_centernet_b200 import decode as D
    from detectron2_centernet_b200 import losses as L
    from detectron2_centernet_b200 import target as T
    with pytest.raises(NotImplementedError):
        D.ctdet_decode(torch.rand(1, 2, 8, 8), torch.rand(1, 2, 8, 8))
    with pytest.raises(NotImplementedError):
        L.FocalLoss([1.0])(torch.rand(1, 2, 8, 8), torch.rand(1, 2, 8, 8))
    with pytest.raises(NotImplementedError):
        T.target_gen(torch.zeros(1, 4), torch.zeros(1, dtype=torch.int64),
                     torch.zeros(2, dtype=torch.int32), 1, 2, 8, 8)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libctdet_b200.so")
    with pytest.raises(RuntimeError, match="no CPU"):
        _lib.load()
