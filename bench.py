#!/usr/bin/env python
"""Headline benchmark: images/sec of a full training step (forward, loss, backward,
gradient all-reduce, optimizer step) through the public ``Stoke`` API.

Default config is BASELINE.json config 2 — ResNet-50, ImageNet-shaped synthetic data,
plain DDP, fp16 AMP via the facade.  Other BASELINE configs: ``--config resnet50_oss_sddp``
(3), ``--config vitb16_fsdp_bf16`` (4), ``--config resnet18_cifar`` (1, also runs on CPU).

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5

``--impl reference`` is the reference arm: the reference (two scripts importing
fairscale / stoke and five modules that are not in its repository) has no installer and
its dependencies are not on the box, so it prints an ``unavailable`` line.
``--impl torch`` runs the torch-native stand-in (torch DDP / ZeRO / FSDP + GradScaler +
torchvision models) that BASELINE.md §3 names; it is labelled a stand-in, never the reference.

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (model, per-GPU batch, input shape, classes, parallelism label, dtype)
    "resnet50_ddp_amp": dict(model="resnet50", batch=256, shape=(3, 224, 224), classes=1000, fp16="amp",
                             oss=False, sddp=False, fsdp=False, accum=1, clip=None, par="dp", dtype="fp16"),
    "resnet50_oss_sddp": dict(model="resnet50", batch=128, shape=(3, 224, 224), classes=1000, fp16="amp",
                              oss=True, sddp=True, fsdp=False, accum=2, clip=0.1, par="zero2", dtype="fp16"),
    "vitb16_fsdp_bf16": dict(model="vit_b_16", batch=128, shape=(3, 224, 224), classes=1000, fp16="bf16",
                             oss=False, sddp=False, fsdp=True, accum=1, clip=None, par="fsdp", dtype="bf16"),
    "resnet18_cifar": dict(model="resnet18", batch=128, shape=(3, 32, 32), classes=10, fp16=None,
                           oss=True, sddp=False, fsdp=False, accum=1, clip=None, par="zero1", dtype="fp32"),
}
ADAMW = {"lr": 1e-3, "betas": (0.9, 0.99), "eps": 1e-8, "weight_decay": 1e-4}  # Stoke-DDP.py:226-235


def parse() -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch"])
    ap.add_argument("--config", default="resnet50_ddp_amp", choices=sorted(CONFIGS))
    ap.add_argument("--batch-size", type=int, default=0, help="per-GPU batch (0: the config's default)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="disable CUDA-graph capture of the train step")
    ap.add_argument("--cpu", action="store_true", help="CPU/gloo plumbing run (config 1)")
    return ap.parse_args()


def maybe_relaunch(args: argparse.Namespace) -> None:
    """``python bench.py --gpus N`` without a launcher: re-exec under torch.distributed.run."""
    if args.gpus > 1 and "RANK" not in os.environ and args.impl != "reference":
        from pytorch_distributedtraining_b200.parallel.comm import find_free_port
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", find_free_port(), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))


def build_ours(cfg: dict, args, world: int):
    import torch
    from pytorch_distributedtraining_b200 import (AMPConfig, B200Config, ClipGradNormConfig, DDPConfig,
                                                  FairscaleFSDPConfig, FairscaleOSSConfig, Stoke, StokeOptimizer)
    from pytorch_distributedtraining_b200 import models
    from pytorch_distributedtraining_b200.ops.loss import cross_entropy

    kw = {"num_classes": cfg["classes"]}
    if cfg["model"] == "resnet18" and cfg["shape"][1] <= 64:
        kw["small_input"] = True
    model = getattr(models, cfg["model"])(**kw)
    gpu = not args.cpu
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    configs = [AMPConfig(init_scale=2.0 ** 14), DDPConfig(local_rank=local_rank, backend="nccl" if gpu else "gloo"),
               FairscaleOSSConfig(broadcast_fp16=False), FairscaleFSDPConfig(mixed_precision=cfg["fp16"] is not None),
               B200Config(cuda_graph=gpu and not args.no_graph)]
    stoke = Stoke(
        model=model, verbose=(int(os.environ.get("RANK", "0")) == 0),
        optimizer=StokeOptimizer(optimizer=torch.optim.AdamW, optimizer_kwargs=dict(ADAMW)),
        loss=cross_entropy, batch_size_per_device=args.batch_size or cfg["batch"], gpu=gpu,
        fp16=cfg["fp16"] if gpu else None, distributed="ddp" if (world > 1 or cfg["oss"] or cfg["fsdp"]) else None,
        fairscale_oss=cfg["oss"], fairscale_sddp=cfg["sddp"], fairscale_fsdp=cfg["fsdp"],
        grad_accum_steps=cfg["accum"], configs=configs,
        grad_clip=None if cfg["clip"] is None else ClipGradNormConfig(max_norm=cfg["clip"], norm_type=2.0))
    stoke.model_access.train()

    def step(x, y):
        return stoke.train_step(x, y)     # public API: forward + loss + backward + (optimizer step)

    return stoke, step, stoke.device


def build_torch(cfg: dict, args, world: int):
    """Torch-native stand-in: torchvision model, torch DDP / ZeRO / FSDP,
    torch.amp autocast + GradScaler, channels_last."""
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    import torchvision

    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    gpu = not args.cpu
    device = torch.device("cuda", local_rank) if gpu else torch.device("cpu")
    if gpu:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl" if gpu else "gloo", device_id=device if gpu else None)
    model = getattr(torchvision.models, cfg["model"])(num_classes=cfg["classes"]).to(device)
    if cfg["model"].startswith("resnet") and gpu:
        model = model.to(memory_format=torch.channels_last)
    amp_dtype = {"amp": torch.float16, "bf16": torch.bfloat16, None: None}[cfg["fp16"] if gpu else None]
    accum = cfg["accum"]
    if cfg["fsdp"] and world > 1:
        from torch.distributed.fsdp import FullyShardedDataParallel as FSDP, MixedPrecision
        from torch.distributed.fsdp.wrap import ModuleWrapPolicy
        from torchvision.models.vision_transformer import EncoderBlock
        mp = MixedPrecision(param_dtype=amp_dtype, reduce_dtype=amp_dtype, buffer_dtype=amp_dtype)
        model = FSDP(model, auto_wrap_policy=ModuleWrapPolicy({EncoderBlock}), mixed_precision=mp,
                     device_id=device, use_orig_params=False)
        opt = torch.optim.AdamW(model.parameters(), fused=True, **ADAMW)
    else:
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank] if gpu else None,
                                                              gradient_as_bucket_view=True)
        if cfg["oss"] and world > 1:
            from torch.distributed.optim import ZeroRedundancyOptimizer
            opt = ZeroRedundancyOptimizer(model.parameters(), optimizer_class=torch.optim.AdamW, **ADAMW)
        else:
            opt = torch.optim.AdamW(model.parameters(), fused=gpu, **ADAMW)
    scaler = torch.amp.GradScaler("cuda", init_scale=2.0 ** 14, enabled=amp_dtype == torch.float16)
    model.train()
    state = {"i": 0}
    import contextlib

    def step(x, y):
        state["i"] += 1
        boundary = state["i"] % accum == 0
        ctx = contextlib.nullcontext() if (boundary or world == 1 or not hasattr(model, "no_sync")) else model.no_sync()
        with ctx:
            with torch.autocast("cuda", dtype=amp_dtype, enabled=amp_dtype is not None):
                loss = F.cross_entropy(model(x), y) / accum
            scaler.scale(loss).backward()
        if boundary:
            if cfg["clip"] is not None:
                scaler.unscale_(opt)
                torch.nn.utils.clip_grad_norm_(model.parameters(), cfg["clip"])
            scaler.step(opt)
            scaler.update()
            opt.zero_grad(set_to_none=True)
        return loss

    return None, step, device


def main() -> None:
    args = parse()
    if args.impl == "reference":
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference is two scripts with no setup.py/pyproject (pip: 'not installable'); its "
                          "fairscale/stoke/wandb deps and 5 local modules are absent and there is no network"}))
        return
    maybe_relaunch(args)
    import torch
    import torch.distributed as dist

    cfg = dict(CONFIGS[args.config])
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if args.cpu and world == 1 and args.gpus > 1:
        world = 1
    build = build_ours if args.impl == "ours" else build_torch
    stoke, step, device = build(cfg, args, world)
    is_cuda = device.type == "cuda"
    B = args.batch_size or cfg["batch"]
    accum = cfg["accum"]

    # ---- synthetic data of the named shape (random-init weights, no datasets on the box)
    g = torch.Generator().manual_seed(1234 + rank)
    n_pool = 2
    host = [(torch.randn((B, *cfg["shape"]), generator=g), torch.randint(0, cfg["classes"], (B,), generator=g))
            for _ in range(n_pool)]
    cl = is_cuda and cfg["model"].startswith("resnet")
    if is_cuda:
        from pytorch_distributedtraining_b200.data import pin_channels_last
        # pinned host batches; for the CNNs stored NHWC on the host (decoded-image order), so the H2D copy
        # lands in the channels-last layout directly
        host = [((pin_channels_last(x) if cl else x.pin_memory()), y.pin_memory()) for x, y in host]

    def to_dev(x, y):
        x = x.to(device, non_blocking=True)
        if cl:
            x = x.contiguous(memory_format=torch.channels_last)
        return x, y.to(device, non_blocking=True)

    dev_pool = [to_dev(x, y) for x, y in host]

    def sync():
        if is_cuda:
            torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[device.index]) if is_cuda else dist.barrier()

    from pytorch_distributedtraining_b200.ops import _lib
    from pytorch_distributedtraining_b200.utils.timing import ClockSampler, max_over_ranks

    # one timed "step" = one optimizer step = `accum` micro-batches
    def run_device_steps(n):
        for i in range(n * accum):
            x, y = dev_pool[i % n_pool]
            step(x, y)

    run_device_steps(max(args.warmup, 3))
    sync(); barrier(); sync()
    sampler = ClockSampler(gpu_index=device.index or 0).start() if (is_cuda and rank == 0) else None
    launches0 = _lib.LAUNCHES
    if is_cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    run_device_steps(args.steps)
    if is_cuda:
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.perf_counter() - t0) * 1e3
    sync(); barrier(); sync()
    launches = _lib.LAUNCHES - launches0
    clocks = sampler.stop() if sampler is not None else None
    ms = max_over_ranks(ms, device if is_cuda else None)
    images = B * accum * world * args.steps
    value = images / (ms / 1e3)

    # ---- end to end: pinned host batch -> H2D every step (prefetched one step ahead on a
    #      copy stream, as StokeDataLoader does) and a D2H read of the loss every step
    e2e = None
    if is_cuda and not args.no_e2e:
        from pytorch_distributedtraining_b200.data import DevicePrefetcher

        def host_batches(n):
            for i in range(n):
                yield host[i % n_pool]

        loss_host = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
        loss_evt = [torch.cuda.Event() for _ in range(2)]

        def run_e2e(n):
            # every step's loss is copied device -> pinned host and READ on the host; the read of step i happens
            # after step i+1 has been enqueued (what a training loop that logs its loss does), so the GPU never
            # idles on the host round trip
            last, pending = 0.0, None
            for i, (x, y) in enumerate(DevicePrefetcher(host_batches(n * accum), device, channels_last=cl)):
                loss = step(x, y)
                slot = i & 1
                loss_host[slot].copy_(loss.detach().float(), non_blocking=True)      # D2H of this step's result
                loss_evt[slot].record()
                if pending is not None:
                    loss_evt[pending].synchronize()
                    last = float(loss_host[pending])
                pending = slot
            if pending is not None:
                loss_evt[pending].synchronize()
                last = float(loss_host[pending])
            return last

        run_e2e(2)
        sync(); barrier(); sync()
        e0.record()
        run_e2e(args.steps)
        e1.record(); e1.synchronize()
        ms_e = max_over_ranks(e0.elapsed_time(e1), device)
        sync(); barrier()
        x0, y0 = host[0]
        e2e = {"value": images / (ms_e / 1e3), "unit": "images/sec",
               "h2d_bytes_per_step": int((x0.numel() * x0.element_size() + y0.numel() * y0.element_size()) * accum),
               "d2h_bytes_per_step": 4 * accum, "ms_per_step": ms_e / args.steps}

    if rank == 0:
        out = {
            "metric": "images/sec", "value": value, "unit": "images/sec", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": cfg["dtype"] if is_cuda else "fp32",
            "data": "synthetic (random images of the named shape, random-init weights)",
            "impl": {"ours": "ours", "torch": "torch-native stand-in (not the reference)"}[args.impl],
            "config": {"model": cfg["model"], "global_batch": B * accum * world, "per_gpu_batch": B,
                       "grad_accum": accum, "input": list(cfg["shape"]), "name": args.config,
                       "parallelism": f"{cfg['par']}{world}", "optimizer": "AdamW",
                       "l2": "per-step working set (inputs + activations, several GB) exceeds the 126 MB L2; "
                             "no explicit flush"},
            "gpu_launches": launches,
            "e2e": e2e,
        }
        if clocks is not None:
            out["clocks"] = clocks
        if is_cuda:
            out["peak_mem_gb"] = round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)
        print(json.dumps(out))
    # teardown: CUDA graphs first (they may hold collective kernels), then the process group; a
    # watchdog makes sure a stuck teardown can never hold the GPUs after the result is printed
    sys.stdout.flush()
    import threading
    threading.Timer(20.0, lambda: os._exit(0)).start()
    if stoke is not None:
        stoke.reset_graphs()
    sync()
    if world > 1 and dist.is_initialized():
        barrier()
        dist.destroy_process_group()
    os._exit(0)


if __name__ == "__main__":
    main()
