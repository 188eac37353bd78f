"""Times ctdet_loss_fwd_bwd (B = 32, 80 x 128 x 128) per tuning-knob setting of the TMA-staged sweep.
The knobs are read once per process, so every setting runs in a child process.
  python scripts/loss_tma_sweep.py            -> table on stdout
  python scripts/loss_tma_sweep.py --child    -> one measurement with the current environment"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child():
    import numpy as np
    import torch
    from detectron2_centernet_b200 import target as T, losses as L, synthetic as S
    dev = "cuda:0"
    B, C, H, W = int(os.environ.get("SWEEP_B", "32")), 80, 128, 128
    bl, cl = S.make_batch_objects(1236, B, 512)
    boxes, classes, offsets = S.pack_objects(bl, cl)
    tb, tc, to = [torch.from_numpy(x).to(dev) for x in (boxes, classes, offsets)]
    g = torch.Generator(device=dev).manual_seed(1)
    hm = torch.randn((B, C, H, W), device=dev, generator=g) * 1.5 - 2.19
    wh = torch.rand((B, 2, H, W), device=dev, generator=g) * 64
    reg = torch.rand((B, 2, H, W), device=dev, generator=g)
    gt = T.target_gen(tb, tc, to, B, C, H, W)
    alpha = L.alpha_tensor([1.0], C, dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def run():
        return L._call(hm, wh, reg, gt["hm"], gt["reg_mask"], gt["ind"], gt["wh"], gt["reg"], alpha, gt["num_pos"],
                       None, 2, (1.0, 0.1, 1.0), True)
    for _ in range(3):
        out = run()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream()
    with torch.cuda.stream(cap):
        run()
        cap.synchronize()
        with torch.cuda.graph(graph, stream=cap):
            run()
    torch.cuda.synchronize()
    for _ in range(3):
        graph.replay()
    ts = []
    for _ in range(30):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); graph.replay(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    losses = [float(v) for v in out[0].cpu()]
    losses.append(float(out[1].double().abs().sum()))        # checksum of d_hm: identical across the variants
    t = float(np.median(ts))
    print("RESULT %.2f us (min %.2f)  %.0f GB/s  losses %s" % (t, min(ts), B * C * H * W * 12 / t / 1e3, losses), flush=True)


def main():
    settings = [{"CTDET_LOSS_TMA": "0"}]
    for q, s in ((2, 4), (2, 3), (2, 6), (1, 8), (1, 4), (4, 2), (4, 3), (2, 8), (2, 2)):
        settings.append({"CTDET_LOSS_TMA": "1", "CTDET_LOSS_TMA_Q": str(q), "CTDET_LOSS_TMA_STAGES": str(s)})
    if len(sys.argv) > 1 and sys.argv[1] == "--only":
        settings = [dict(kv.split("=") for kv in a.split(",")) for a in sys.argv[2:]]
    for st in settings:
        env = dict(os.environ)
        env.update(st)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"], env=env, capture_output=True, text=True)
        line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")]
        print(st, line[0] if line else "FAILED: " + r.stderr[-600:], flush=True)


if __name__ == "__main__":
    child() if (len(sys.argv) > 1 and sys.argv[1] == "--child") else main()
