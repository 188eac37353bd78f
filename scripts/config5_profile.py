"""Kernel-level breakdown of one config-5 step (neck + head + decode, batch 32) with the torch profiler."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from torch.profiler import profile, ProfilerActivity
from detectron2_centernet_b200 import neck as N
from detectron2_centernet_b200 import decode as D
dev = torch.device("cuda", 0)
torch.manual_seed(1239)
neck = N.DLA34Neck().to(dev).eval()
head = N.CtdetHead(64, 80, 256).to(dev).eval()
for m in neck.modules():
    if hasattr(m, "conv_offset_mask"):
        torch.nn.init.normal_(m.conv_offset_mask.weight, std=1.0 / (9 * m.conv_offset_mask.in_channels) ** 0.5)
B = 32
feats = [torch.randn(B, c, 512 >> i, 512 >> i, device=dev) for i, c in enumerate((16, 32, 64, 128, 256, 512))]
def step():
    y = neck(feats)[-1]
    out = head(y)
    return D.ctdet_decode_batch(out["hm"], out["wh"], out["reg"], down_ratio=4, K=100, logits=True)
with torch.no_grad():
    for _ in range(3): step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3): step()
        torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
