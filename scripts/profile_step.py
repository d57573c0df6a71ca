"""Kernel-time breakdown of one training step with torch.profiler (CUPTI).  Shares only —
numbers under a profiler are never bench values."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import argparse
import torch
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--impl", default="ours")
ap.add_argument("--config", default="resnet50_ddp_amp")
ap.add_argument("--batch-size", type=int, default=0)
ap.add_argument("--out", default="gpurun_out/prof.txt")
a = ap.parse_args()
args = argparse.Namespace(cpu=False, batch_size=a.batch_size, impl=a.impl, no_graph=True, sync_bn=False)
cfg = dict(bench.CONFIGS[a.config])
build = bench.build_ours if a.impl == "ours" else bench.build_torch
stoke, step, device = build(cfg, args, 1)
B = a.batch_size or cfg["batch"]
x = torch.randn(B, *cfg["shape"], device=device)
if cfg["model"].startswith("resnet"):
    x = x.contiguous(memory_format=torch.channels_last)
if cfg.get("target_scale"):
    sc = cfg["target_scale"]
    x = torch.rand(B, *cfg["shape"], device=device)
    y = torch.rand(B, cfg["shape"][0], cfg["shape"][1] * sc, cfg["shape"][2] * sc, device=device)
else:
    y = torch.randint(0, cfg["classes"], (B,), device=device)
for _ in range(6):
    step(x, y)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step(x, y)
    torch.cuda.synchronize()
tab = prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=90)
os.makedirs(os.path.dirname(a.out), exist_ok=True)
open(a.out, "w").write(tab)
print(tab[-6000:])
