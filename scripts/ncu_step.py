"""ncu launch target: ONE eager ResNet-50 training step (batch 256, fp16 AMP, the headline config) between
cudaProfilerStart / Stop, after warm-up.  Used with `--profile-from-start off` and a short metric list to put the
tensor-pipe / DRAM shares beside every kernel of the step (profiles/resnet50_step_kernels.md)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import argparse
import torch
import bench

args = argparse.Namespace(cpu=False, batch_size=0, impl="ours", no_graph=True, sync_bn=False)
cfg = dict(bench.CONFIGS["resnet50_ddp_amp"])
stoke, step, device = bench.build_ours(cfg, args, 1)
B = cfg["batch"]
x = torch.randn(B, *cfg["shape"], device=device).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, cfg["classes"], (B,), device=device)
for _ in range(4):
    step(x, y)
torch.cuda.synchronize()
torch.cuda.profiler.start()
step(x, y)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
