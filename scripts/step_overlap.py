"""Graph topologies of the device-resident step (target-gen -> loss, and the decode that only reads the head outputs):
which overlap hides the decode's latency-bound ends.  B = 32, L2 flushed (memset + read) between replays."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from detectron2_centernet_b200 import target as T, losses as L, decode as D, synthetic as S
dev = "cuda:0"
B, C, H, W, K = 32, 80, 128, 128, 100
bl, cl = S.make_batch_objects(1236, B, 512)
boxes, classes, offsets = S.pack_objects(bl, cl)
tb, tc, to = [torch.from_numpy(x).to(dev) for x in (boxes, classes, offsets)]
hm_, wh_, reg_ = S.make_head_outputs(1242, B, C, H, W, bl)
hm, wh, reg = [torch.from_numpy(x).to(dev) for x in (hm_, wh_, reg_)]
gt = T.target_gen(tb, tc, to, B, C, H, W)
alpha = L.alpha_tensor([1.0], C, dev)
fl_w = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
fl_r = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
tgen = lambda: T.target_gen(tb, tc, to, B, C, H, W, out=gt)
loss = lambda: L._call(hm, wh, reg, gt["hm"], gt["reg_mask"], gt["ind"], gt["wh"], gt["reg"], alpha, gt["num_pos"], None, 2, (1.0, 0.1, 1.0), True)
dec = lambda: D.ctdet_decode_batch(hm, wh, reg, down_ratio=4, K=K, logits=True)
cur = torch.cuda.current_stream


def serial(side):
    tgen(); loss(); dec()

def decode_first_serial(side):
    dec(); tgen(); loss()

def fork_at_start(side):                    # bench.py's two-branch graph
    side.wait_stream(cur())
    with torch.cuda.stream(side):
        dec()
    tgen(); loss()
    cur().wait_stream(side)

def decode_beside_target_gen_only(side):    # the loss waits for the decode as well
    side.wait_stream(cur())
    with torch.cuda.stream(side):
        dec()
    tgen()
    cur().wait_stream(side)
    loss()

def decode_beside_loss(side):               # the decode starts with the loss
    tgen()
    side.wait_stream(cur())
    with torch.cuda.stream(side):
        dec()
    loss()
    cur().wait_stream(side)

for prio in (0, -1):
    for fn in (serial, decode_first_serial, fork_at_start, decode_beside_target_gen_only, decode_beside_loss):
        main = torch.cuda.Stream()
        side = torch.cuda.Stream(priority=prio)
        with torch.cuda.stream(main):
            fn(side); main.synchronize(); side.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=main):
                fn(side)
        torch.cuda.synchronize()
        for _ in range(3):
            g.replay()
        evs = []
        for _ in range(40):
            fl_w.zero_(); fl_r.sum()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); g.replay(); e.record(); evs.append((s, e))
        torch.cuda.synchronize()
        ts = [s.elapsed_time(e) * 1e3 for s, e in evs]
        print("side priority %2d  %-32s mean %.1f us  median %.1f" % (prio, fn.__name__, float(np.mean(ts)), float(np.median(ts))), flush=True)
