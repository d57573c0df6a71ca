#!/usr/bin/env python
"""The reference's first entry point (``Fairscale-DDP.py``) on this framework: ``mp.spawn`` N
processes, ``init_process_group``, paired-image dataset split + ``DistributedSampler`` +
``DataLoader``, a sub-pixel SR CNN, ``OSS(AdamW)`` + ``ShardedDDP``, a plain training loop
(Fairscale-DDP.py:20-133).  Differences are only the reference's bugs: ``os`` is
imported, ``num_replicas`` follows the world size instead of a hard-coded 4, tensors are moved to
the device when one is available, and the sampler's epoch advances.

    python examples/fairscale_ddp.py --world-size 2 --epochs 1            # CPU / gloo, synthetic data
    python examples/fairscale_ddp.py --world-size 2 --gpu                 # one GPU per rank, NCCL
    python examples/fairscale_ddp.py --input-dir LR/ --target-dir HR/     # real paired image folders
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn as nn
from torch.optim import AdamW
from torch.utils.data import DataLoader

from pytorch_distributedtraining_b200 import (OSS, CustomDataset, DistributedSampler, ShardedDDP, SyntheticPairedDataset,
                                              spawn_training)
from pytorch_distributedtraining_b200.models import Net


def train(rank: int, world_size: int, epochs: int, args):
    gpu = args.gpu and torch.cuda.is_available()
    if gpu:
        torch.cuda.set_device(rank)
    dist.init_process_group(backend="nccl" if gpu else "gloo", init_method="env://", rank=rank, world_size=world_size)
    device = torch.device("cuda", rank) if gpu else torch.device("cpu")
    print("===> Loading datasets")
    if args.input_dir:
        full_dataset = CustomDataset(args.input_dir, args.target_dir)
    else:
        full_dataset = SyntheticPairedDataset(args.samples, lr_size=32, scale=2)
    train_size = int(0.99 * len(full_dataset))
    train_dataset, val_dataset = torch.utils.data.random_split(
        full_dataset, [train_size, len(full_dataset) - train_size], generator=torch.Generator().manual_seed(0))
    train_sampler = DistributedSampler(train_dataset, num_replicas=world_size, rank=rank)
    val_sampler = DistributedSampler(val_dataset, num_replicas=world_size, rank=rank)
    training_dataloader = DataLoader(dataset=train_dataset, num_workers=args.workers, batch_size=args.batch_size,
                                     drop_last=True, shuffle=False, pin_memory=gpu, sampler=train_sampler)
    val_dataloader = DataLoader(dataset=val_dataset, num_workers=0, batch_size=args.batch_size, shuffle=False,
                                sampler=val_sampler)
    print("Length of Training dataset - ", len(train_dataset))

    print("===> Building model")
    model = Net(upscale_factor=2).to(device)
    loss_fn = nn.MSELoss()
    optimizer = OSS(params=model.parameters(), optim=AdamW, lr=1e-3, betas=(0.9, 0.99), eps=1e-8, weight_decay=1e-4)
    model = ShardedDDP(model, optimizer)     # reduces each gradient to the rank that owns it

    model.train()
    loss = torch.zeros(())
    for e in range(epochs):
        for iteration, batch in enumerate(training_dataloader, 1):
            model.zero_grad()
            outputs = model(batch[0].to(device, non_blocking=True))
            loss = loss_fn(outputs, batch[1].to(device, non_blocking=True))
            loss.backward()
            optimizer.step()
            if iteration % 25 == 0 and rank == 0:
                print(loss.item())
        if rank == 0:
            print("For Epoch {}, loss: {:.4f}".format(e, float(loss)))
    dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--world-size", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch-size", type=int, default=8)
    ap.add_argument("--samples", type=int, default=256)
    ap.add_argument("--workers", type=int, default=0)
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--input-dir", default="")
    ap.add_argument("--target-dir", default="")
    a = ap.parse_args()
    spawn_training(train, a.world_size, args=(a.epochs, a))
