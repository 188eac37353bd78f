"""Times ctdet_head_decode_fused (C ABI, preallocated outputs, CUDA graph replay, L2 flushed) against cuDNN conv + ctdet_decode.
usage: python scripts/head_decode_time.py [B] [relu]   (relu: class-concentrated logits -> spill list)"""
import sys, ctypes, torch
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from detectron2_centernet_b200 import _lib
dev = 'cuda:0'
B, Cf, C, H, W, K = int(sys.argv[1]) if len(sys.argv) > 1 else 32, 256, 80, 128, 128, 100
lib = _lib.load()
g = torch.Generator(device='cpu').manual_seed(1)
feat = torch.randn(B, Cf, H, W, generator=g)
if len(sys.argv) > 2: feat = torch.relu(feat)
feat = feat.to(dev)
w = torch.randn(C, Cf, generator=g)
w = (w / w.norm(dim=1, keepdim=True) * 1.0).to(dev)
b = torch.full((C,), -2.19, device=dev)
wh = torch.rand(B, 2, H, W, device=dev) * 20
reg = torch.rand(B, 2, H, W, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = (torch.empty(B, K, 4, device=dev), torch.empty(B, K, device=dev), torch.empty(B, K, dtype=torch.int32, device=dev),
       torch.empty(B, K, dtype=torch.int64, device=dev))
n1 = lib.ctdet_head_decode_fused_workspace_bytes(B, Cf, C, H, W, K)
ws1 = torch.zeros(n1, dtype=torch.uint8, device=dev)
n2 = lib.ctdet_decode_workspace_bytes(B, C, H, W, K)
ws2 = torch.zeros(n2, dtype=torch.uint8, device=dev)
w4 = w[:, :, None, None].contiguous()
logits = torch.empty(B, C, H, W, device=dev)
def fused():
    rc = lib.ctdet_head_decode_fused(_lib.ptr(feat), _lib.ptr(w), _lib.ptr(b), _lib.ptr(wh), _lib.ptr(reg), B, Cf, C, H, W, K, 4.0, 0,
                                     *[_lib.ptr(o) for o in out], None, _lib.ptr(ws1), ctypes.c_size_t(n1), _lib.stream_ptr(dev))
    assert rc == 0
def dec():
    rc = lib.ctdet_decode(_lib.ptr(logits), _lib.ptr(wh), _lib.ptr(reg), B, C, H, W, K, 4.0, 1, *[_lib.ptr(o) for o in out],
                          _lib.ptr(ws2), ctypes.c_size_t(n2), _lib.stream_ptr(dev))
    assert rc == 0
def conv():
    torch.nn.functional.conv2d(feat, w4, b, ) if False else torch.ops.aten.convolution.out(feat, w4, b, [1, 1], [0, 0], [1, 1], False, [0, 0], 1, out=logits)
def unfused():
    conv(); dec()
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(gr, stream=s):
            fn()
    ts = []
    for _ in range(n):
        flush.zero_(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
print('B', B, 'fused us %.1f' % t(fused), ' conv %.1f' % t(conv), ' decode %.1f' % t(dec), ' conv+decode %.1f' % t(unfused))
