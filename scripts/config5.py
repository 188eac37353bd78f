"""BASELINE config 5: DLA-34 IDAUp/DLAUp neck (16 DCNv2 layers on the tcgen05 kernels) + full ctdet head + top-100
decode, 512 images sharded by image over N GPUs (micro-batch 32 per GPU per step), eval mode, synthetic base-network
features (the DLA base is outside the hot path: SURVEY 8 a14).  One process per GPU:

    python scripts/config5.py                                  # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/config5.py

CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks; prints one JSON line.
--offsets random|zero: a random-init DCN has a zero-initialised offset conv (offset 0, mask 0.5: integer taps);
`random` re-initialises it so that the offsets spread over about +-2 pixels and is the headline."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import torch.distributed as dist
from detectron2_centernet_b200 import neck as N
from detectron2_centernet_b200 import decode as D
from detectron2_centernet_b200.dist import shard_range


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=512)
    ap.add_argument("--micro", type=int, default=32)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--offsets", default="random", choices=["random", "zero"])
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lo, hi = shard_range(args.images, rank, world)
    n_mine = hi - lo
    torch.manual_seed(1239 + rank)
    neck = N.DLA34Neck().to(dev).eval()
    head = N.CtdetHead(64, 80, 256).to(dev).eval()
    if args.offsets == "random":
        for m in neck.modules():
            if hasattr(m, "conv_offset_mask"):
                torch.nn.init.normal_(m.conv_offset_mask.weight, std=1.0 / (3.0 * (9 * m.conv_offset_mask.in_channels) ** 0.5) * 3.0)
    B = args.micro
    chans, S = (16, 32, 64, 128, 256, 512), 512
    feats = [torch.randn(B, c, S >> i, S >> i, device=dev) for i, c in enumerate(chans)]

    def step():
        y = neck(feats)[-1]
        out = head(y)
        return D.ctdet_decode_batch(out["hm"], out["wh"], out["reg"], down_ratio=4, K=100, logits=True)

    def segment_times():
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record(); y = neck(feats)[-1]
        ev[1].record(); out = head(y)
        ev[2].record(); D.ctdet_decode_batch(out["hm"], out["wh"], out["reg"], down_ratio=4, K=100, logits=True)
        ev[3].record(); torch.cuda.synchronize()
        return [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]

    with torch.no_grad():
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        seg = segment_times()
        steps = (n_mine + B - 1) // B
        best = None
        for _ in range(args.reps):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(steps):
                step()
            e.record(); torch.cuda.synchronize()
            ms = torch.tensor([s.elapsed_time(e)], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            best = ms.item() if best is None else min(best, ms.item())
    if rank == 0:
        gflop = 14.19          # main DCN contraction per image
        print(json.dumps({"config": "cfg5: DLA-34 neck (16 DCNv2) + ctdet head + top-100 decode, eval, %d images over %d GPU(s), micro-batch %d"
                          % (args.images, world, B), "offsets": args.offsets, "n_gpus": world, "ms_total": round(best, 3),
                          "images_per_s": round(args.images / best * 1e3, 1), "steps_per_gpu": steps,
                          "per_step_ms": {"neck": round(seg[0], 3), "head": round(seg[1], 3), "decode": round(seg[2], 3)},
                          "dcn_tflops_if_neck_were_all_dcn": round(gflop * B / seg[0], 1)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
