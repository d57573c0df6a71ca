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
    # the reference's own workload (Stoke-DDP.py:206-254): SwinIR-S x2 (embed_dim 60, 6 heads, window 8), 64x64 patches,
    # batch 18 / GPU, OSS + ShardedDDP, accumulation 2, clip 0.1, fp16 owner broadcast, perceptual loss.  The reference
    # leaves AMP off (fp32); here the compute dtype is bf16 (the B200 default), stated in `dtype`.
    "swinir_s_oss_sddp": dict(model="swinir_s", batch=18, shape=(3, 64, 64), classes=0, fp16="bf16",
                              oss=True, sddp=True, fsdp=False, accum=2, clip=0.1, par="zero2", dtype="bf16",
                              loss="feat_loss", target_scale=2, broadcast_fp16=True),
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
    ap.add_argument("--sync-bn", action="store_true", help="convert BatchNorm to SyncBatchNorm (Stoke-DDP.py:192) in both arms")
    ap.add_argument("--no-standin", action="store_true", help="skip the same-session torch stand-in arm")
    ap.add_argument("--no-exposed", action="store_true", help="skip the exposed-communication measurement (N > 1)")
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

    if cfg["model"] == "swinir_s":
        model = models.SwinIR(upscale=2, in_chans=3, img_size=64, window_size=8, img_range=1., depths=[6, 6, 6, 6],
                              embed_dim=60, num_heads=[6, 6, 6, 6], mlp_ratio=2, upsampler="pixelshuffledirect",
                              resi_connection="1conv")
    else:
        kw = {"num_classes": cfg["classes"]}
        if cfg["model"] == "resnet18" and cfg["shape"][1] <= 64:
            kw["small_input"] = True
        model = getattr(models, cfg["model"])(**kw)
    loss_fn = cross_entropy
    if cfg.get("loss") == "feat_loss":
        from pytorch_distributedtraining_b200.models.perceptual import feat_loss as loss_fn
    gpu = not args.cpu
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    configs = [AMPConfig(init_scale=2.0 ** 14),
               DDPConfig(local_rank=local_rank, backend="nccl" if gpu else "gloo", convert_to_sync_batch_norm=args.sync_bn),
               FairscaleOSSConfig(broadcast_fp16=bool(cfg.get("broadcast_fp16", False))), FairscaleFSDPConfig(mixed_precision=cfg["fp16"] is not None),
               B200Config(cuda_graph=gpu and not args.no_graph, strict=gpu)]
    stoke = Stoke(
        model=model, verbose=(int(os.environ.get("RANK", "0")) == 0),
        optimizer=StokeOptimizer(optimizer=torch.optim.AdamW, optimizer_kwargs=dict(ADAMW)),
        loss=loss_fn, batch_size_per_device=args.batch_size or cfg["batch"], gpu=gpu,
        fp16=cfg["fp16"] if gpu else None, distributed="ddp" if (world > 1 or cfg["oss"] or cfg["fsdp"]) else None,
        fairscale_oss=cfg["oss"], fairscale_sddp=cfg["sddp"], fairscale_fsdp=cfg["fsdp"],
        grad_accum_steps=cfg["accum"], configs=configs,
        grad_clip=None if cfg["clip"] is None else ClipGradNormConfig(max_norm=cfg["clip"], norm_type=2.0))
    stoke.model_access.train()

    def step(x, y):
        return stoke.train_step(x, y)     # public API: forward + loss + backward + (optimizer step)

    return stoke, step, stoke.device


def build_torch(cfg: dict, args, world: int, capturable: bool = False):
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
    if cfg["model"] == "swinir_s":
        # no torch-native SwinIR exists: the stand-in runs THIS repository's module definition with every custom
        # kernel switched off (stock cuBLAS / cuDNN / SDPA underneath), wrapped in torch DDP + ZeroRedundancyOptimizer
        os.environ["B200DP_DISABLE"] = "conv,gemm,bn,ln,ce,pool,mse,convstats,attn_layout,mlp,dgrad_s2,attn,pad"
        from pytorch_distributedtraining_b200.ops import dispatch
        for f in os.environ["B200DP_DISABLE"].split(","):
            dispatch.disable(f)
        from pytorch_distributedtraining_b200 import models as our_models
        model = our_models.SwinIR(upscale=2, in_chans=3, img_size=64, window_size=8, img_range=1., depths=[6, 6, 6, 6],
                                  embed_dim=60, num_heads=[6, 6, 6, 6], mlp_ratio=2, upsampler="pixelshuffledirect",
                                  resi_connection="1conv").to(device)
    else:
        model = getattr(torchvision.models, cfg["model"])(num_classes=cfg["classes"]).to(device)
    if cfg["model"].startswith("resnet") and gpu:
        model = model.to(memory_format=torch.channels_last)
    if getattr(args, "sync_bn", False) and world > 1:
        model = torch.nn.SyncBatchNorm.convert_sync_batchnorm(model)
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
            opt = torch.optim.AdamW(model.parameters(), fused=gpu, capturable=capturable and gpu, **ADAMW)
    scaler = torch.amp.GradScaler("cuda", init_scale=2.0 ** 14, enabled=amp_dtype == torch.float16)
    t_loss = None
    if cfg.get("loss") == "feat_loss":
        from pytorch_distributedtraining_b200.models.perceptual import feat_loss as t_loss
    model.train()
    state = {"i": 0}
    import contextlib

    def step(x, y):
        state["i"] += 1
        boundary = state["i"] % accum == 0
        ctx = contextlib.nullcontext() if (boundary or world == 1 or not hasattr(model, "no_sync")) else model.no_sync()
        with ctx:
            with torch.autocast("cuda", dtype=amp_dtype, enabled=amp_dtype is not None):
                loss = (t_loss(model(x), y) if t_loss is not None else F.cross_entropy(model(x), y)) / accum
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


# fwd+bwd model FLOPs per image (2 x MACs x 3: forward + input gradient + weight gradient), for the roofline line
MODEL_GFLOP_PER_IMAGE = {"resnet50": 24.6, "resnet18": 10.9, "vit_b_16": 105.4}


def graph_torch_step(step, x, y, device):
    """Whole-step CUDA-graph capture of the stand-in (single process only), so the comparison with the
    graph-captured arm isolates kernels instead of launch overhead.  Returns a replay callable."""
    import torch
    sx, sy = x.clone(), y.clone()
    side = torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(side):
        for _ in range(3):
            step(sx, sy)
    torch.cuda.current_stream(device).wait_stream(side)
    torch.cuda.synchronize(device)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out = step(sx, sy)

    def replay(x, y):
        sx.copy_(x, non_blocking=True)
        sy.copy_(y, non_blocking=True)
        g.replay()
        return out

    return replay


def measure(step, args, cfg, device, world, rank, host, dev_pool, cl, want_e2e, count_launches, graphs_pending=None):
    """Device-timed steps (CUDA events, barrier + synchronize on both sides, max over ranks), then the same
    through the host path: pinned batch -> H2D every step, loss -> D2H every step."""
    import torch
    import torch.distributed as dist
    from pytorch_distributedtraining_b200.ops import _lib
    from pytorch_distributedtraining_b200.utils.timing import ClockSampler, max_over_ranks
    is_cuda = device.type == "cuda"
    B = args.batch_size or cfg["batch"]
    accum, n_pool = cfg["accum"], len(dev_pool)

    def sync():
        if is_cuda:
            torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[device.index]) if is_cuda else dist.barrier()

    # one timed "step" = one optimizer step = `accum` micro-batches
    def run_device_steps(n):
        loss = None
        for i in range(n * accum):
            x, y = dev_pool[i % n_pool]
            loss = step(x, y)
        return loss

    run_device_steps(max(args.warmup, 3))
    extra = 0
    while graphs_pending is not None and graphs_pending() and extra < 8:
        # every accumulation phase needs three eager calls before its graph is captured: keep warming up (untimed)
        # until all of them replay, so no capture lands inside the timed region
        run_device_steps(1)
        extra += 1
    sync(); barrier(); sync()
    sampler = ClockSampler(gpu_index=device.index or 0).start() if (is_cuda and rank == 0) else None
    launches0 = _lib.LAUNCHES
    if is_cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    last = run_device_steps(args.steps)
    if is_cuda:
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.perf_counter() - t0) * 1e3
    sync(); barrier(); sync()
    launches = _lib.LAUNCHES - launches0 if count_launches else 0
    clocks = sampler.stop() if sampler is not None else None
    ms = max_over_ranks(ms, device if is_cuda else None)
    images = B * accum * world * args.steps
    res = {"value": images / (ms / 1e3), "ms_per_step": ms / args.steps, "launches": launches, "clocks": clocks,
           "warmup_done": max(args.warmup, 3) + extra,
           "final_loss": float(last.detach().float()) if last is not None else None, "e2e": None}

    # ---- end to end: pinned host batch -> H2D every step (prefetched one step ahead on a
    #      copy stream, as StokeDataLoader does) and a D2H read of the loss every step
    if is_cuda and want_e2e:
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
        res["e2e"] = {"value": images / (ms_e / 1e3), "unit": "images/sec",
                      "h2d_bytes_per_step": int((x0.numel() * x0.element_size() + y0.numel() * y0.element_size()) * accum),
                      "d2h_bytes_per_step": 4 * accum, "ms_per_step": ms_e / args.steps}
    return res


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
    graph_standin = args.impl == "torch" and not args.cpu and world == 1 and not args.no_graph and cfg["accum"] == 1
    if args.impl == "ours":
        stoke, step, device = build_ours(cfg, args, world)
    else:
        stoke, step, device = build_torch(cfg, args, world, capturable=graph_standin)
    is_cuda = device.type == "cuda"
    B = args.batch_size or cfg["batch"]
    accum = cfg["accum"]
    if is_cuda and args.impl == "ours":
        from pytorch_distributedtraining_b200.ops import require_library
        require_library()            # a GPU run without the kernel library is an error, never a silent fallback

    # ---- synthetic data of the named shape (random-init weights, no datasets on the box)
    g = torch.Generator().manual_seed(1234 + rank)
    n_pool = 2
    if cfg.get("target_scale"):          # super-resolution: low-res input, high-res target
        c, hh, ww = cfg["shape"]
        sc = cfg["target_scale"]
        host = [(torch.rand((B, c, hh, ww), generator=g), torch.rand((B, c, hh * sc, ww * sc), generator=g)) for _ in range(n_pool)]
    else:
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
    standin_graphed = False
    if graph_standin:
        step = graph_torch_step(step, *dev_pool[0], device)
        standin_graphed = True

    pending = None
    if stoke is not None and is_cuda and not args.no_graph:
        def pending():
            g = stoke._graphs
            return len(g) < accum or any(e.get("graph") is None for e in g.values())
    res = measure(step, args, cfg, device, world, rank, host, dev_pool, cl, not args.no_e2e, args.impl == "ours", pending)
    native = stoke.native_status if stoke is not None else None

    # ---- same session, same box: the torch-native stand-in (NOT the reference, which cannot be installed)
    standin = None
    if args.impl == "ours" and is_cuda and not args.no_standin:
        try:
            graphed = world == 1 and accum == 1 and not args.no_graph
            _, t_step, _ = build_torch(cfg, args, world, capturable=graphed)
            if graphed:
                t_step = graph_torch_step(t_step, *dev_pool[0], device)
            t = measure(t_step, args, cfg, device, world, rank, host, dev_pool, cl, not args.no_e2e, False)
            standin = {"what": "torch DDP/ZeRO/FSDP + torchvision + torch.amp + cuDNN/cuBLAS/NCCL, same process group; "
                               "a stand-in, not the reference", "cuda_graph": graphed, "value": t["value"],
                       "ms_per_step": t["ms_per_step"], "e2e_value": t["e2e"]["value"] if t["e2e"] else None,
                       "ratio": res["value"] / t["value"],
                       "e2e_ratio": (res["e2e"]["value"] / t["e2e"]["value"]) if (t["e2e"] and res["e2e"]) else None}
            del t_step
        except Exception as e:  # the stand-in must never take the product's number down with it
            standin = {"error": f"{type(e).__name__}: {str(e)[:200]}"}

    # ---- exposed communication: the same captured step with every peer-memory collective muted.  Run last:
    #      with the collectives off the replicas drift apart, so nothing may be measured after this.
    exposed = None
    if args.impl == "ours" and is_cuda and world > 1 and not args.no_exposed:
        from pytorch_distributedtraining_b200.ops import collectives as C
        stoke.reset_graphs()
        C.MUTE = True
        try:
            muted = measure(step, args, cfg, device, world, rank, host, dev_pool, cl, False, False, pending)
            diff = res["ms_per_step"] - muted["ms_per_step"]
            # a muted step that comes out SLOWER is not a measurement (with the collectives off, replicas diverge and the
            # optimizer state goes stale): report nothing rather than a negative number
            exposed = {"exposed_comm_ms": diff if diff > -0.05 * res["ms_per_step"] else None, "ms_per_step_comm_muted": muted["ms_per_step"],
                       "how": "same whole-step CUDA graph re-captured with all peer-memory collective launches skipped"}
        finally:
            C.MUTE = False

    if rank == 0:
        gflop = MODEL_GFLOP_PER_IMAGE.get(cfg["model"])
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        out = {
            "metric": "images/sec", "value": res["value"], "unit": "images/sec", "n_gpus": world, "steps": args.steps,
            "warmup": res["warmup_done"], "ms_per_step": res["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": cfg["dtype"] if is_cuda else "fp32",
            "data": "synthetic (random images of the named shape, random-init weights)",
            "impl": {"ours": "ours", "torch": "torch-native stand-in (not the reference)"}[args.impl],
            "config": {"model": cfg["model"], "global_batch": B * accum * world, "per_gpu_batch": B,
                       "grad_accum": accum, "input": list(cfg["shape"]), "name": args.config,
                       "parallelism": f"{cfg['par']}{world}", "optimizer": "AdamW", "sync_bn": bool(args.sync_bn),
                       "l2": "per-step working set (inputs + activations, several GB) exceeds the 126 MB L2; "
                             "no explicit flush"},
            "gpu_launches": res["launches"],
            "e2e": res["e2e"],
            "final_loss": res["final_loss"],
            "native": native,
            "standin": standin,
        }
        if args.impl == "torch":
            out["cuda_graph"] = standin_graphed
        if exposed is not None:
            out.update(exposed)
        if gflop is not None and is_cuda:
            tflops = res["value"] * gflop / 1e3 / world
            out["roofline"] = {"model_tflops_per_gpu": tflops,
                               "frac_of_bf16_sustained": tflops / peaks["bf16_tflops_sustained"] if peaks.get("bf16_tflops_sustained") else None,
                               "peaks": "MEASURED_PEAKS.json" if peaks else "absent"}
        if res["clocks"] is not None:
            out["clocks"] = res["clocks"]
        if is_cuda:
            out["peak_mem_gb"] = round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)
        print(json.dumps(out))

    def sync():
        if is_cuda:
            torch.cuda.synchronize()

    # teardown: CUDA graphs first (they may hold collective kernels), then the process group, then a NORMAL
    # interpreter exit so exit hooks run.  The daemon timer only fires if teardown is stuck for minutes.
    sys.stdout.flush()
    import threading
    guard = threading.Timer(180.0, lambda: (sys.stderr.write("[bench] teardown stuck for 180 s; forcing exit\n"),
                                            sys.stderr.flush(), os._exit(3)))
    guard.daemon = True
    guard.start()
    if stoke is not None:
        stoke.reset_graphs()
    sync()
    if dist.is_initialized():            # also the one-rank group the ZeRO / FSDP configs create on a single GPU
        if world > 1:
            dist.barrier(device_ids=[device.index]) if is_cuda else dist.barrier()
        dist.destroy_process_group()
    guard.cancel()


if __name__ == "__main__":
    main()
