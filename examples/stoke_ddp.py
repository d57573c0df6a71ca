#!/usr/bin/env python
"""The reference's second entry point (``Stoke-DDP.py``) on this framework, same structure:
argparse flags, AMP / DDP / OSS config objects, SwinIR-S x2, perceptual loss, a ``Stoke`` object
with OSS + ShardedDDP + gradient accumulation + norm clipping, Stoke DataLoaders with
DistributedSamplers, OneCycleLR, train / validate / checkpoint per epoch, run logging
(Stoke-DDP.py:47-339).  Launch it the way the reference is launched:

    torchrun --nproc-per-node 4 examples/stoke_ddp.py --batchSize 18 --nEpochs 2 --gpu
    python examples/stoke_ddp.py --nEpochs 1 --samples 64                 # single process, CPU, synthetic

Deliberate fixes: ``--lr`` is honoured, the LR scheduler steps once per optimizer step,
validation metrics are averaged across ranks, logging is rank-0-only and works offline
(JSONL under ``runs/`` when W&B is not importable / reachable), ``--resume`` restores a checkpoint.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.optim as optim
from torch.optim import AdamW

from pytorch_distributedtraining_b200 import (AMPConfig, ClipGradNormConfig, CustomDataset, DDPConfig, DistributedOptions,
                                              DistributedSampler, FairscaleOSSConfig, FP16Options, Stoke, StokeOptimizer,
                                              SyntheticPairedDataset)
from pytorch_distributedtraining_b200.launch import env_rank_info
from pytorch_distributedtraining_b200.models import SwinIR, feat_loss
from pytorch_distributedtraining_b200.utils import metrics
from pytorch_distributedtraining_b200.utils.logging import RunLogger, train_log, val_log


def train(train_dataloader, stoke_model: Stoke, scheduler1, epoch: int, logger: RunLogger):
    example_ct, batch_ct, sum_loss = 0, 0, 0.0
    stoke_model.print_on_devices(f"Starting Epoch {epoch + 1}")
    stoke_model.model_access.train()
    for idx, (inputs, targets) in enumerate(train_dataloader):
        outputs = stoke_model.model(inputs)
        train_loss = stoke_model.loss(outputs, targets)
        if (idx + 1) % 10 == 0:
            stoke_model.print_ema_loss(prepend_msg=f"Step {idx + 1} -- EMA Loss")
        stoke_model.backward(loss=train_loss)
        steps_before = stoke_model.optimizer_steps
        stoke_model.step()
        if stoke_model.optimizer_steps != steps_before:
            scheduler1.step()                       # once per optimizer step, not per micro-batch
        sum_loss += stoke_model.detach_and_sync_loss(loss=train_loss)
        example_ct += len(inputs)
        batch_ct += 1
        if ((batch_ct + 1) % 50) == 0:
            train_log(logger, float(train_loss), example_ct, epoch)
    return sum_loss / max(len(train_dataloader), 1)


def validate(val_dataloader, stoke_model: Stoke, epoch: int, logger: RunLogger):
    stoke_model.model_access.eval()
    val_loss = torch.zeros((), device=stoke_model.device)
    mae = torch.zeros((), device=stoke_model.device)
    psnr = torch.zeros((), device=stoke_model.device)
    example_ct = 0
    with torch.no_grad():
        for inputs, targets in val_dataloader:
            example_ct += len(inputs)
            outputs = stoke_model.model(inputs)
            val_loss += stoke_model.loss(outputs, targets)
            mae += metrics.mae(outputs, targets)
            psnr += metrics.psnr(outputs, targets)
    n = max(len(val_dataloader), 1)
    val_avg_loss = float(metrics.reduce_mean(val_loss / n))
    avg_mae, avg_psnr = float(metrics.reduce_mean(mae / n)), float(metrics.reduce_mean(psnr / n))
    val_log(logger, val_avg_loss, avg_mae, avg_psnr, example_ct, epoch)
    stoke_model.print_on_devices(msg=f"Current Average Validation Loss: {val_avg_loss}, PSNR : {avg_psnr}")
    return val_avg_loss


def save_checkpoint(stoke_model: Stoke, epoch, train_loss, val_loss):
    path, tag = stoke_model.save(path="checkpoint/", name="model_{}_{:.2f}_{:.2f}".format(epoch, train_loss, val_loss),
                                 extras={"epoch": epoch})
    stoke_model.print_on_devices("Checkpoint saved after epoch {} -> {}/{}".format(epoch, path, tag))
    return tag


def main():
    parser = argparse.ArgumentParser(description="B200 sharded data-parallel super-resolution training")
    parser.add_argument("--projectName", default="Stoke-4K-2X-DDP", type=str)
    parser.add_argument("--batchSize", type=int, default=18)
    parser.add_argument("--nEpochs", type=int, default=10)
    parser.add_argument("--start-epoch", default=0, type=int)
    parser.add_argument("--resume", default="", type=str, help="checkpoint tag under checkpoint/ to resume from")
    parser.add_argument("--lr", type=float, default=0.001)
    parser.add_argument("--weight_decay", "--wd", default=1e-4, type=float)
    parser.add_argument("--grad_clip", type=float, default=0.1)
    parser.add_argument("--local_rank", default=-1, type=int)
    parser.add_argument("--threads", type=int, default=0)
    parser.add_argument("--inputDir", type=str, default="")
    parser.add_argument("--targetDir", type=str, default="")
    parser.add_argument("--samples", type=int, default=512, help="synthetic samples when no --inputDir is given")
    parser.add_argument("--gpu", action="store_true")
    parser.add_argument("--amp", action="store_true")
    opt = parser.parse_args()
    rank, local_rank, world = env_rank_info()
    epochs = opt.nEpochs

    amp_config = AMPConfig(init_scale=2.0 ** 14)
    ddp_config = DDPConfig(local_rank=local_rank, convert_to_sync_batch_norm=True,
                           backend="nccl" if opt.gpu else "gloo")
    oss_config = FairscaleOSSConfig(broadcast_fp16=True)

    print("===> Building model")
    model = SwinIR(upscale=2, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6, 6, 6, 6], embed_dim=60,
                   num_heads=[6, 6, 6, 6], mlp_ratio=2, upsampler="pixelshuffledirect", resi_connection="1conv")
    model_path = "model_zoo/002_lightweightSR_DIV2K_s64w8_SwinIR-S_x2.pth"
    if os.path.exists(model_path):
        pretrained = torch.load(model_path, map_location="cpu")
        model.load_state_dict(pretrained["params"] if "params" in pretrained else pretrained, strict=True)

    optimizer = StokeOptimizer(optimizer=AdamW, optimizer_kwargs={"lr": opt.lr, "betas": (0.9, 0.99), "eps": 1e-8,
                                                                   "weight_decay": opt.weight_decay})
    distributed = world > 1
    stoke_model = Stoke(
        model=model, verbose=True, optimizer=optimizer, loss=feat_loss, batch_size_per_device=opt.batchSize,
        gpu=opt.gpu, fp16=FP16Options.amp.value if (opt.amp and opt.gpu) else None,
        distributed=DistributedOptions.ddp.value if distributed else None,
        fairscale_oss=distributed, fairscale_sddp=distributed, grad_accum_steps=2,
        configs=[amp_config, ddp_config, oss_config],
        grad_clip=ClipGradNormConfig(max_norm=opt.grad_clip, norm_type=2.0))

    print("===> Loading datasets")
    full_dataset = CustomDataset(opt.inputDir, opt.targetDir) if opt.inputDir else SyntheticPairedDataset(opt.samples, 64, 2)
    train_size = int(0.9 * len(full_dataset))
    train_dataset, val_dataset = torch.utils.data.random_split(
        full_dataset, [train_size, len(full_dataset) - train_size], generator=torch.Generator().manual_seed(0))
    train_sampler = DistributedSampler(dataset=train_dataset, num_replicas=stoke_model.world_size, rank=stoke_model.rank)
    val_sampler = DistributedSampler(val_dataset, num_replicas=stoke_model.world_size, rank=stoke_model.rank)
    train_dataloader = stoke_model.DataLoader(dataset=train_dataset, sampler=train_sampler, num_workers=opt.threads,
                                              multiprocessing_context="spawn" if opt.threads else None)
    val_dataloader = stoke_model.DataLoader(dataset=val_dataset, sampler=val_sampler, num_workers=0)

    steps_per_epoch = max(len(train_dataloader) // stoke_model.grad_accum, 1)
    scheduler1 = optim.lr_scheduler.OneCycleLR(stoke_model.optimizer, max_lr=0.01, pct_start=0.9,
                                               steps_per_epoch=steps_per_epoch, epochs=epochs)
    scheduler2 = optim.lr_scheduler.ReduceLROnPlateau(stoke_model.optimizer, mode="min", factor=0.2, patience=2, min_lr=5e-5)
    logger = RunLogger(project=opt.projectName, rank=stoke_model.rank,
                       config=dict(epochs=opt.nEpochs, batch_size=opt.batchSize, learning_rate=opt.lr,
                                   dataset="synthetic" if not opt.inputDir else opt.inputDir, architecture="SwinIR-S-x2"))
    start = opt.start_epoch
    if opt.resume:
        extras = stoke_model.load("checkpoint/", opt.resume)
        start = (extras or {}).get("epoch", -1) + 1

    print("===> Training")
    for epoch in range(start, epochs):
        train_loss = train(train_dataloader, stoke_model, scheduler1, epoch, logger)
        val_loss = validate(val_dataloader, stoke_model, epoch, logger)
        save_checkpoint(stoke_model, epoch, train_loss, val_loss)
        stoke_model.print_on_devices("--------Train Loss after Epoch {} - {} --------".format(epoch, train_loss))
        stoke_model.print_on_devices("--------Val Loss after Epoch {} - {} --------".format(epoch, val_loss))
    logger.finish()


if __name__ == "__main__":
    main()
