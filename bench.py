#!/usr/bin/env python
"""bench.py -- ctdet heatmap hot path throughput (BASELINE.json metric: images/sec @512^2).

Workload (config.workload = BASELINE.json configs[1] + the decode of the metric): one "step" is
one pass of the hot path over a batch of 32 synthetic COCO-shaped images at 512^2 (80x128x128
heatmap): Gaussian target generation -> fused focal/RegL1 loss forward+backward -> top-100 decode.

  python bench.py [--gpus N] [--steps K] [--warmup W]           our CUDA path (one rank per GPU)
  python bench.py --impl reference ...                          the reference's CPU path (oracle port)

One JSON line on stdout (rank 0).  `value` = images/s with inputs resident in HBM; `e2e` = the same
through the public API from pinned HOST buffers, H2D/D2H inside the timed region; `roofline` =
the dominant kernel (fused loss) against the measured HBM peak; `cpu_baseline` = the oracle port
timed on this box's host cores on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from detectron2_centernet_b200 import synthetic as S   # noqa: E402

B_PER_GPU, C, H, W, K, INPUT = 32, 80, 128, 128, 100, 512
METRIC = "ctdet target-gen + focal/RegL1 loss fwd+bwd + top-100 decode, images/sec @512^2"
WORKLOAD = ("ctdet DLA-34 training-step heads, batch=32/GPU @512^2 (80x128x128 heatmap, <=40 objects/image): "
            "gaussian targets + fused focal/RegL1 loss fwd+bwd + top-100 decode with 3x3 peak NMS")
BYTES_PER_CELL = {"target_gen": 4, "loss": 12, "decode": 4, "fused_target_loss": 8}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def tensor_peak():
    """Sustained dense bf16 TFLOP/s (the DCN stack is timed inside a long loop)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1389.0))), "measured (sustained)"
    return 1389.0, "fallback"


# the 16 DeformConvV2 layers of the DLA-34 neck at 512^2: (Cin, Cout, H=W, count)
DCN_LAYERS = [(512, 256, 16, 1), (256, 256, 32, 1), (256, 128, 32, 2), (128, 128, 64, 2), (128, 64, 64, 4),
              (64, 64, 128, 5), (256, 64, 32, 1)]


def dcn_stack(dev, B, reps=5):
    """BASELINE.json configs[4] (kernel part): forward and backward of the 16 DCNv2 layers of the
    DLA-34 IDAUp/DLAUp neck, random offsets in (-2, 2), sigmoid masks; CUDA events, device-resident."""
    import math
    from detectron2_centernet_b200 import deform_conv as DC
    g = torch.Generator(device=dev).manual_seed(0)
    fwd_ms = bwd_ms = flops = 0.0
    for Cin, Cout, S_, cnt in DCN_LAYERS:
        x = torch.randn((B, Cin, S_, S_), device=dev, generator=g)
        off = (torch.rand((B, 18, S_, S_), device=dev, generator=g) - 0.5) * 4
        m = torch.sigmoid(torch.randn((B, 9, S_, S_), device=dev, generator=g))
        w = (torch.rand((Cout, Cin, 3, 3), device=dev, generator=g) - 0.5) * 2 / math.sqrt(9 * Cin)
        bias = torch.zeros(Cout, device=dev)
        gy = torch.randn((B, Cout, S_, S_), device=dev, generator=g)
        for which in ("fwd", "bwd"):
            if which == "fwd":
                fn = lambda: DC.dcnv2_forward(x, off, m, w, bias, 1, 1, 1)          # noqa: E731
            else:
                fn = lambda: DC.dcnv2_backward(x, off, m, w, gy, 1, 1, 1)           # noqa: E731
            for _ in range(2):
                fn()
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(reps):
                fn()
            e.record()
            torch.cuda.synchronize()
            ms = s.elapsed_time(e) / reps
            if which == "fwd":
                fwd_ms += ms * cnt
            else:
                bwd_ms += ms * cnt
        flops += 2.0 * Cout * Cin * 9 * S_ * S_ * B * cnt
    peak, src = tensor_peak()
    tf = flops / fwd_ms / 1e9
    return {"workload": "16 DCNv2 layers of the DLA-34 neck @512^2, batch %d, fp32 NCHW in/out, bf16 operands" % B,
            "fwd_ms": round(fwd_ms, 3), "fwd_images_per_s": round(B / fwd_ms * 1e3, 1), "fwd_tflops": round(tf, 1),
            "bwd_ms": round(bwd_ms, 3), "bwd_tflops": round(2 * flops / bwd_ms / 1e9, 1),
            "gflop_per_image_fwd": round(flops / B / 1e9, 2),
            "roofline": {"bound": "tensor", "achieved": round(tf, 1), "peak": peak, "unit": "TFLOP/s",
                         "frac": round(tf / peak, 4), "peak_source": src,
                         "note": "the sampler is L1-bandwidth-bound (8 B of L1 reads per im2col element), see DESIGN.md section 5"}}


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled through NVML every 5 ms DURING the timed regions."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.sm, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ready = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                     "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                     "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                     "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            while not self._stop_evt.is_set():
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
                self.ready.set()
                time.sleep(0.005)
        except Exception as e:                                  # NVML missing: report it, do not fake
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)
            self.ready.set()

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ----------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist
    from detectron2_centernet_b200 import decode as D
    from detectron2_centernet_b200 import losses as L
    from detectron2_centernet_b200 import target as T
    from detectron2_centernet_b200 import _lib

    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _lib.load()
    B = B_PER_GPU                                   # weak scaling: 32 images per GPU per step
    cells = B * C * H * W

    # per-rank synthetic inputs (seed = 1236 + rank), generated on the host like a data loader would
    bl, cl = S.make_batch_objects(1236 + rank, B, INPUT)
    boxes, classes, offsets = S.pack_objects(bl, cl)
    hm_np, wh_np, reg_np = S.make_head_outputs(2236 + rank, B, C, H, W, bl, tie_free=False)
    pin = lambda a: torch.from_numpy(a).pin_memory()            # noqa: E731
    h_boxes, h_classes, h_offsets = pin(boxes), pin(classes), pin(offsets)
    h_hm, h_wh, h_reg = pin(hm_np), pin(wh_np), pin(reg_np)
    d_boxes, d_classes, d_offsets = h_boxes.to(dev), h_classes.to(dev), h_offsets.to(dev)
    d_hm, d_wh, d_reg = h_hm.to(dev), h_wh.to(dev), h_reg.to(dev)
    alpha = L.alpha_tensor([1.0], C, dev)
    weights = (1.0, 0.1, 1.0)
    gt = T.target_gen(d_boxes, d_classes, d_offsets, B, C, H, W)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2
    h_out = torch.empty(3 + B * K * 6, dtype=torch.float32).pin_memory()

    def hot_path(hm, wh, reg, bx, cls_, off):
        """target-gen -> fused loss fwd+bwd -> decode (seed, sweep, tail): 5 kernel launches of ours."""
        T.target_gen(bx, cls_, off, B, C, H, W, out=gt)
        losses, g_hm, g_wh, g_reg = L._call(hm, wh, reg, gt["hm"], gt["reg_mask"], gt["ind"], gt["wh"],
                                            gt["reg"], alpha, gt["num_pos"], None, 2, weights, True)
        det = D.ctdet_decode_batch(hm, wh, reg, down_ratio=4, K=K, logits=True)
        return losses, det

    # kernels launched per step: target_gen + loss_kernel + the decode's launches -- 3 on the streaming path
    # (decode_seed, decode_stream, decode_tail; the default for these 128x128 maps), 2 on the staged path
    # (decode_peaks, decode_tail; CTDET_DECODE_STREAM=0)
    DECODE_LAUNCHES = 2 if os.environ.get("CTDET_DECODE_STREAM", "1") == "0" else 3
    LAUNCHES_PER_STEP = 2 + DECODE_LAUNCHES

    # The device-resident step is captured ONCE into a CUDA graph (5 kernel nodes) and replayed:
    # the kernels take 30-100 us each, so per-call Python/ctypes overhead would otherwise show up
    # as idle gaps on the stream.
    for _ in range(2):
        hot_path(d_hm, d_wh, d_reg, d_boxes, d_classes, d_offsets)          # warm caches / workspaces
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    cap_stream = torch.cuda.Stream()
    with torch.cuda.stream(cap_stream):
        # workspaces are keyed by stream: create them on the capture stream before capturing
        hot_path(d_hm, d_wh, d_reg, d_boxes, d_classes, d_offsets)
        cap_stream.synchronize()
        with torch.cuda.graph(graph, stream=cap_stream):
            graph_out = hot_path(d_hm, d_wh, d_reg, d_boxes, d_classes, d_offsets)
    torch.cuda.synchronize()

    def step_device():
        graph.replay()
        return graph_out

    # Separate variant: targets generated inside the loss kernel, the
    # heatmap target never reaches HBM (8 B/cell on the training side instead of 16).
    slots = {k: v for k, v in gt.items() if k != "hm"}

    def hot_path_fused(hm, wh, reg, bx, cls_, off):
        L._FusedTargetLoss.apply(hm, wh, reg, bx, cls_, off, alpha, None, 2, weights, slots)
        return D.ctdet_decode_batch(hm, wh, reg, down_ratio=4, K=K, logits=True)

    for _ in range(2):
        hot_path_fused(d_hm, d_wh, d_reg, d_boxes, d_classes, d_offsets)
    torch.cuda.synchronize()
    graph_f = torch.cuda.CUDAGraph()
    with torch.cuda.stream(cap_stream):
        hot_path_fused(d_hm, d_wh, d_reg, d_boxes, d_classes, d_offsets)
        cap_stream.synchronize()
        with torch.cuda.graph(graph_f, stream=cap_stream):
            hot_path_fused(d_hm, d_wh, d_reg, d_boxes, d_classes, d_offsets)
    torch.cuda.synchronize()

    def step_e2e():
        hm = h_hm.to(dev, non_blocking=True)
        wh = h_wh.to(dev, non_blocking=True)
        reg = h_reg.to(dev, non_blocking=True)
        bx = h_boxes.to(dev, non_blocking=True)
        cl_ = h_classes.to(dev, non_blocking=True)
        off = h_offsets.to(dev, non_blocking=True)
        losses, det = hot_path(hm, wh, reg, bx, cl_, off)
        packed = torch.cat([losses, det[0].reshape(-1), det[1].reshape(-1), det[2].reshape(-1).float()])
        h_out.copy_(packed, non_blocking=True)
        torch.cuda.current_stream().synchronize()            # the result is on the host
        return h_out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, flush_l2):
        for _ in range(warmup):
            fn()
        barrier()
        evs = []
        for _ in range(steps):
            if flush_l2:
                flush.zero_()                                # evict L2 between timed iterations
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            evs.append((s, e))
        barrier()
        ms = sum(s.elapsed_time(e) for s, e in evs)          # device time of the K steps
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)         # max over ranks
        return float(t.item())

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        sampler.ready.wait(timeout=5)
    total_ms = timed(step_device, args.steps, args.warmup, True)
    fused_ms = timed(graph_f.replay, args.steps, args.warmup, True)

    # dominant kernel alone (fused loss), CUDA events on the launching stream, L2 flushed
    def loss_only():
        L._call(d_hm, d_wh, d_reg, gt["hm"], gt["reg_mask"], gt["ind"], gt["wh"], gt["reg"], alpha,
                gt["num_pos"], None, 2, weights, True)
    per_kernel = {}
    for name, fn in (("loss", loss_only),
                     ("target_gen", lambda: T.target_gen(d_boxes, d_classes, d_offsets, B, C, H, W, out=gt)),
                     ("decode", lambda: D.ctdet_decode_batch(d_hm, d_wh, d_reg, down_ratio=4, K=K, logits=True)),
                     ("fused_target_loss", lambda: L._FusedTargetLoss.apply(d_hm, d_wh, d_reg, d_boxes, d_classes, d_offsets,
                                                                             alpha, None, 2, weights, slots))):
        ms = timed(fn, max(args.steps, 10), 3, True) / max(args.steps, 10)
        per_kernel[name] = {"us": round(ms * 1e3, 2),
                            "GBps": round(cells * BYTES_PER_CELL[name] / (ms * 1e-3) / 1e9, 1)}
    e2e_ms = timed(step_e2e, args.steps, args.warmup, False)
    clocks = sampler.stop() if sampler else None               # covers every timed region above

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    ms_per_step = total_ms / args.steps
    value = world * B / (ms_per_step * 1e-3)
    e2e_value = world * B / (e2e_ms / args.steps * 1e-3)
    h2d = sum(t.numel() * t.element_size() for t in (h_hm, h_wh, h_reg, h_boxes, h_classes, h_offsets))
    ach = per_kernel["loss"]["GBps"]
    out = {
        "metric": METRIC, "value": round(value, 1), "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic (seeded COCO-shaped boxes; logits ~N(-2.19,1.5^2) + planted peaks)",
        "config": {"workload": WORKLOAD, "images_per_gpu": B, "heatmap": [C, H, W], "K": K,
                   "l2": "flushed (256 MB memset) between timed iterations",
                   "launch": "device-resident step replayed from a CUDA graph (%d kernel nodes); e2e uses plain calls" % LAUNCHES_PER_STEP,
                   "sharding": "by image, no data-path collective"},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": "images/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": h_out.numel() * 4,
                "note": "head outputs (hm/wh/reg, 176 MB/step) copied from pinned host memory every step: PCIe-bound"},
        "gpu_launches": LAUNCHES_PER_STEP * args.steps,
        "roofline": {"bound": "hbm", "kernel": "ctdet::loss_kernel (fused focal+RegL1 fwd+bwd)",
                     "achieved": ach, "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": cells * 12, "traffic": None},
        "kernels": per_kernel,
        "fused_variant": {
            "what": "same step with ctdet_target_loss_fused: the Gaussian target is built in shared memory inside the loss "
                    "sweep and never written to HBM (SURVEY.md section 8 f2; 12 B/cell per step instead of 20)",
            "value": round(world * B / (fused_ms / args.steps * 1e-3), 1), "unit": "images/s",
            "ms_per_step": round(fused_ms / args.steps, 4),
            "gpu_launches_per_step": 2 + DECODE_LAUNCHES},   # target_slots + target_loss_fused + decode
    }
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(traffic_file):
        out["roofline"]["traffic"] = json.load(open(traffic_file)).get("loss_kernel_bytes_per_launch")
    if not args.no_dcn:
        out["dcn_stack"] = dcn_stack(dev, B)
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(sample_images=2, reps=2)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------------------------
def cpu_step(bl, cl, hm, wh, reg):
    """The reference's CPU path for the same step (oracle port of gen_heatmap, CenterNet.losses +
    backward, ctdet_decode); test infrastructure used here only as the timed CPU baseline."""
    from oracle import ctdet_oracle as O
    gt = O.gen_heatmap_batch(bl, cl, H, W, C)
    O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), gt)
    O.decode_torch(S.sigmoid_clamp(hm), wh, reg, 4, K)


def cpu_baseline(sample_images=2, reps=2):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    bl, cl = S.make_batch_objects(1236, sample_images, INPUT)
    hm, wh, reg = S.make_head_outputs(2236, sample_images, C, H, W, bl, tie_free=False)
    cpu_step(bl, cl, hm, wh, reg)                                  # warm-up
    t0 = time.perf_counter()
    for _ in range(reps):
        cpu_step(bl, cl, hm, wh, reg)
    dt = (time.perf_counter() - t0) / reps
    return {"value": round(sample_images / dt, 2), "unit": "images/s", "cores": cores, "kind": "port",
            "sample": "%d images of the same workload x %d reps (oracle port of the reference's numpy/torch CPU path)"
                      % (sample_images, reps)}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n = 2                                                          # images per bounded step
    bl, cl = S.make_batch_objects(1236, n, INPUT)
    hm, wh, reg = S.make_head_outputs(2236, n, C, H, W, bl, tie_free=False)
    for _ in range(args.warmup):
        cpu_step(bl, cl, hm, wh, reg)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_step(bl, cl, hm, wh, reg)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    sample = "each step = %d images of the workload on the host cores (oracle port; the reference package itself " \
             "needs fvcore/yacs/pycocotools which are absent)" % n
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(value, 2), "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "images_per_step": n},
        "cpu_baseline": {"value": round(value, 2), "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 2), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dcn", action="store_true", help="skip the DCNv2 stack timing (extra key dcn_stack)")
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
