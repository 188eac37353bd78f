"""Times the 16 DCNv2 layers of the DLA-34 neck at 512^2 on one GPU."""
import sys, math, json
import numpy as np, torch
sys.path.insert(0, '.')
from detectron2_centernet_b200 import deform_conv as DC
dev = 'cuda:0'
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
fp32 = len(sys.argv) > 3 and sys.argv[3] == 'fp32'
# (Cin, Cout, H=W, count)
LAYERS = [(512, 256, 16, 1), (256, 256, 32, 1), (256, 128, 32, 2), (128, 128, 64, 2), (128, 64, 64, 4),
          (64, 64, 128, 5), (256, 64, 32, 1)]
g = torch.Generator(device=dev).manual_seed(0)
tot_ms, tot_fl = 0.0, 0.0
rows = []
for Cin, Cout, S, cnt in LAYERS:
    x = torch.randn((B, Cin, S, S), device=dev, generator=g)
    off = (torch.rand((B, 18, S, S), device=dev, generator=g) - 0.5) * 4
    m = torch.sigmoid(torch.randn((B, 9, S, S), device=dev, generator=g))
    w = (torch.rand((Cout, Cin, 3, 3), device=dev, generator=g) - 0.5) * 2 / math.sqrt(9 * Cin)
    b = torch.zeros(Cout, device=dev)
    for _ in range(2): DC.dcnv2_forward(x, off, m, w, b, 1, 1, 1, fp32_blend=fp32)
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): DC.dcnv2_forward(x, off, m, w, b, 1, 1, 1, fp32_blend=fp32)
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / reps
    fl = 2.0 * Cout * Cin * 9 * S * S * B
    rows.append((Cin, Cout, S, cnt, ms, fl / ms / 1e9))
    tot_ms += ms * cnt; tot_fl += fl * cnt
for r in rows: print('%4d->%3d @%3d^2 x%d : %.3f ms  %.1f TFLOP/s' % r)
if len(sys.argv) > 4 and sys.argv[4] == 'bwd':
    tb = 0.0
    for Cin, Cout, S, cnt in LAYERS:
        x = torch.randn((B, Cin, S, S), device=dev, generator=g)
        off = (torch.rand((B, 18, S, S), device=dev, generator=g) - 0.5) * 4
        m = torch.sigmoid(torch.randn((B, 9, S, S), device=dev, generator=g))
        w = (torch.rand((Cout, Cin, 3, 3), device=dev, generator=g) - 0.5) * 2 / math.sqrt(9 * Cin)
        gy = torch.randn((B, Cout, S, S), device=dev, generator=g)
        res = {}
        for name, kw in (('wgrad', dict(need_input_grads=False)), ('dgrad', dict(need_weight_grad=False)), ('both', {})):
            for _ in range(2): DC.dcnv2_backward(x, off, m, w, gy, 1, 1, 1, **kw)
            torch.cuda.synchronize()
            s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(reps): DC.dcnv2_backward(x, off, m, w, gy, 1, 1, 1, **kw)
            e.record(); torch.cuda.synchronize()
            res[name] = s.elapsed_time(e) / reps
        fl = 2.0 * Cout * Cin * 9 * S * S * B * 2          # dgrad GEMM + wgrad GEMM
        print('bwd %4d->%3d @%3d^2 x%d : wgrad %.3f  dgrad %.3f  both %.3f ms  (%.1f TFLOP/s)' %
              (Cin, Cout, S, cnt, res['wgrad'], res['dgrad'], res['both'], fl / res['both'] / 1e9))
        tb += res['both'] * cnt
    print(json.dumps({"B": B, "dcn_stack_bwd_ms": round(tb, 3), "fwd_plus_bwd_images_per_s": round(B / (tb + tot_ms) * 1e3, 1)}))
print(json.dumps({"B": B, "dcn_stack_ms": round(tot_ms, 3), "images_per_s": round(B / tot_ms * 1e3, 1),
                  "tflops": round(tot_fl / tot_ms / 1e9, 1), "gflop_per_image": round(tot_fl / B / 1e9, 2), "fp32_blend": fp32}))
