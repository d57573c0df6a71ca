"""Semantic tests of the sharded engines on CPU/gloo, world_size 2.

Oracle: parameters after N steps == single-process training on the
concatenated global batch with a plain torch optimizer.
"""
import copy

import pytest
import torch
import torch.distributed as dist
import torch.nn as nn

from pytorch_distributedtraining_b200.parallel import (OSS, DistributedDataParallel, FullyShardedDataParallel,
                                                        ShardedDataParallel, ShardedGradScaler, SyncBatchNorm,
                                                        auto_wrap, convert_sync_batchnorm)
from pytorch_distributedtraining_b200.utils.testing import spawn

ADAMW = dict(lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=1e-4)


def make_model(bn: bool = False) -> nn.Module:
    torch.manual_seed(7)
    layers = [nn.Linear(12, 32), nn.ReLU(), nn.Linear(32, 24)]
    if bn:
        layers += [nn.BatchNorm1d(24)]
    layers += [nn.ReLU(), nn.Linear(24, 5)]
    return nn.Sequential(*layers)


def make_data(world: int, steps: int, per_rank: int = 4):
    g = torch.Generator().manual_seed(99)
    xs = torch.randn(steps, world * per_rank, 12, generator=g)
    ys = torch.randn(steps, world * per_rank, 5, generator=g)
    return xs, ys


def reference_params(world: int, steps: int, bn=False, accum=1, clip=None):
    model = make_model(bn)
    opt = torch.optim.AdamW(model.parameters(), **ADAMW)
    xs, ys = make_data(world, steps)
    for s in range(steps):
        loss = nn.functional.mse_loss(model(xs[s]), ys[s]) / accum
        loss.backward()
        if (s + 1) % accum == 0:
            if clip is not None:
                torch.nn.utils.clip_grad_norm_(model.parameters(), clip)
            opt.step()
            opt.zero_grad()
    return [p.detach().clone() for p in model.parameters()], opt.state_dict()


def _shard(t, rank, world):
    n = t.shape[0] // world
    return t[rank * n:(rank + 1) * n]


def _assert_close(ps, ref, tol=1e-5):
    for a, b in zip(ps, ref):
        torch.testing.assert_close(a, b, rtol=tol, atol=tol)


# ----------------------------------------------------------------------------- OSS + ShardedDDP
def _w_oss_sddp(rank, world, steps, accum, clip, bucket):
    model = make_model()
    if rank != 0:  # start-up sync must repair this
        with torch.no_grad():
            for p in model.parameters():
                p.add_(1.0)
    opt = OSS(model.parameters(), optim=torch.optim.AdamW, **ADAMW)
    sddp = ShardedDataParallel(model, opt, reduce_buffer_size=bucket)
    xs, ys = make_data(world, steps)
    import contextlib
    for s in range(steps):
        boundary = (s + 1) % accum == 0
        ctx = contextlib.nullcontext() if boundary else sddp.no_sync()
        with ctx:
            loss = nn.functional.mse_loss(sddp(_shard(xs[s], rank, world)), _shard(ys[s], rank, world)) / accum
            loss.backward()
        if boundary:
            # owner holds the reduced gradient, others dropped theirs
            for p in model.parameters():
                owner = opt.param_to_rank[p]
                assert (p.grad is not None) == (owner == rank), "gradient sharding violated"
            if clip is not None:
                opt.clip_grad_norm(clip)
            opt.step()
            sddp.zero_grad()
    ref, ref_sd = reference_params(world, steps, accum=accum, clip=clip)
    _assert_close(list(model.parameters()), ref)
    # checkpoint parity: consolidated state loads into a plain torch optimizer
    opt.consolidate_state_dict(recipient_rank=0)
    if rank == 0:
        sd = opt.state_dict()
        assert set(sd.keys()) == {"state", "param_groups"}
        assert sorted(sd["state"].keys()) == sorted(ref_sd["state"].keys())
        plain = torch.optim.AdamW(make_model().parameters(), **ADAMW)
        plain.load_state_dict(sd)
        for k in ref_sd["state"]:
            torch.testing.assert_close(sd["state"][k]["exp_avg"], ref_sd["state"][k]["exp_avg"], rtol=1e-5, atol=1e-6)
    # and a full state dict loads back into a sharded optimizer
    full = dist.broadcast_object_list
    box = [opt.state_dict() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    opt2 = OSS(make_model().parameters(), optim=torch.optim.AdamW, **ADAMW)
    opt2.load_state_dict(box[0])
    assert len(opt2.optim.state) == len(opt.local_params())


@pytest.mark.parametrize("accum,clip,bucket", [(1, None, 2 ** 23), (2, 0.1, 2 ** 23), (1, 0.5, 0), (2, None, 600)])
def test_oss_sharded_ddp_matches_single_process(accum, clip, bucket):
    spawn(_w_oss_sddp, 2, (4, accum, clip, bucket))


def test_oss_partition_is_balanced_and_deterministic():
    model = make_model()
    opt = OSS(model.parameters(), optim=torch.optim.SGD, lr=0.1)
    assert opt.world_size == 1 and len(opt.partition_parameters()) == 1
    # emulate world 3 partition without a process group
    opt.world_size = 3
    opt._clear_cache()
    part = opt.partition_parameters()
    sizes = [sum(p.numel() for g in r for p in g["params"]) for r in part]
    assert sum(sizes) == sum(p.numel() for p in model.parameters())
    assert max(sizes) - min(sizes) <= max(p.numel() for p in model.parameters())
    opt._clear_cache()
    assert [[id(p) for g in r for p in g["params"]] for r in opt.partition_parameters()] == \
           [[id(p) for g in r for p in g["params"]] for r in part]


def _w_oss_fp16_wire(rank, world):
    model = make_model()
    opt = OSS(model.parameters(), optim=torch.optim.AdamW, broadcast_fp16=True, **ADAMW)
    sddp = ShardedDataParallel(model, opt)
    xs, ys = make_data(world, 3)
    for s in range(3):
        nn.functional.mse_loss(sddp(_shard(xs[s], rank, world)), _shard(ys[s], rank, world)).backward()
        opt.step()
        sddp.zero_grad()
    # replicas identical, values fp16-representable, close to the fp32 run
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    other = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(other, flat)
    assert torch.equal(other[0], other[1])
    assert torch.equal(flat, flat.half().float())
    ref, _ = reference_params(world, 3)
    _assert_close(list(model.parameters()), ref, tol=2e-3)
    # the owner kept fp32 masters
    assert all(m.dtype == torch.float32 for m in opt._master.values()) and len(opt._master) > 0


def test_oss_broadcast_fp16_keeps_replicas_identical():
    spawn(_w_oss_fp16_wire, 2)


# ----------------------------------------------------------------------------- DDP
def _w_ddp(rank, world, accum, hook):
    model = make_model()
    ddp = DistributedDataParallel(model, bucket_cap_mb=0.002)
    assert len(ddp.buckets) > 1
    if hook:
        from pytorch_distributedtraining_b200.parallel import bf16_compress_hook
        ddp.register_comm_hook(None, bf16_compress_hook)
    opt = torch.optim.AdamW(model.parameters(), **ADAMW)
    xs, ys = make_data(world, 4)
    import contextlib
    for s in range(4):
        boundary = (s + 1) % accum == 0
        with (contextlib.nullcontext() if boundary else ddp.no_sync()):
            (nn.functional.mse_loss(ddp(_shard(xs[s], rank, world)), _shard(ys[s], rank, world)) / accum).backward()
        if boundary:
            opt.step()
            ddp.zero_grad()
    ref, _ = reference_params(world, 4, accum=accum)
    _assert_close(list(model.parameters()), ref, tol=5e-2 if hook else 1e-5)


@pytest.mark.parametrize("accum,hook", [(1, False), (2, False), (1, True)])
def test_ddp_matches_single_process(accum, hook):
    spawn(_w_ddp, 2, (accum, hook))


# ----------------------------------------------------------------------------- FSDP
def _w_fsdp(rank, world, nested, reshard, accum):
    model = make_model()
    if nested:
        model = auto_wrap(model, policy=lambda m, n: n >= 100, reshard_after_forward=reshard)
    fsdp = FullyShardedDataParallel(model, reshard_after_forward=reshard)
    opt = torch.optim.AdamW(fsdp.parameters(), **ADAMW)
    n_local = sum(p.numel() for p in fsdp.parameters())
    n_total = sum(p.numel() for p in make_model().parameters())
    assert n_local < n_total * 0.75, "parameters are not sharded"
    xs, ys = make_data(world, 4)
    import contextlib
    for s in range(4):
        boundary = (s + 1) % accum == 0
        with (contextlib.nullcontext() if boundary else fsdp.no_sync()):
            (nn.functional.mse_loss(fsdp(_shard(xs[s], rank, world)), _shard(ys[s], rank, world)) / accum).backward()
        if boundary:
            opt.step()
            opt.zero_grad()
    ref, _ = reference_params(world, 4, accum=accum)
    sd = fsdp.state_dict()
    names = [k for k, _ in make_model().named_parameters()]
    assert list(sd.keys()) == names or set(sd.keys()) == set(names)
    _assert_close([sd[k] for k in names], ref)
    # round trip through a full state dict
    fsdp2 = FullyShardedDataParallel(make_model(), reshard_after_forward=reshard)
    fsdp2.load_state_dict(sd)
    sd2 = fsdp2.state_dict()
    for k in names:
        assert torch.equal(sd[k], sd2[k])
    # clip + summon
    total = fsdp.clip_grad_norm_(1.0) if any(p.grad is not None for p in fsdp.parameters()) else None
    with fsdp.summon_full_params():
        assert fsdp.module[0].weight.shape == (32, 12) if not nested else True
    full_os = fsdp.gather_full_optim_state_dict(opt)
    assert set(full_os["state"].keys()) == set(names)
    assert full_os["state"]["0.weight"]["exp_avg"].shape == (32, 12)


@pytest.mark.parametrize("nested,reshard,accum", [(False, True, 1), (True, True, 1), (True, False, 2), (False, True, 2)])
def test_fsdp_matches_single_process(nested, reshard, accum):
    spawn(_w_fsdp, 2, (nested, reshard, accum))


# ----------------------------------------------------------------------------- SyncBN
def _w_syncbn(rank, world):
    # SGD, not Adam: the bias in front of a BatchNorm has a true gradient of zero and Adam
    # would amplify its rounding noise to +-lr, which makes the comparison ill-conditioned.
    model = convert_sync_batchnorm(make_model(bn=True))
    assert any(isinstance(m, SyncBatchNorm) for m in model.modules())
    ddp = DistributedDataParallel(model)
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
    ref_model = make_model(bn=True)
    opt_ref = torch.optim.SGD(ref_model.parameters(), lr=0.1, momentum=0.9)
    xs, ys = make_data(world, 3)
    for s in range(3):
        nn.functional.mse_loss(ddp(_shard(xs[s], rank, world)), _shard(ys[s], rank, world)).backward()
        opt.step()
        ddp.zero_grad()
        nn.functional.mse_loss(ref_model(xs[s]), ys[s]).backward()
        opt_ref.step()
        opt_ref.zero_grad()
    _assert_close(list(model.parameters()), list(ref_model.parameters()), tol=2e-5)
    # running stats equal the global-batch stats
    torch.testing.assert_close(model[3].running_var, ref_model[3].running_var, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(model[3].running_mean, ref_model[3].running_mean, rtol=1e-4, atol=1e-5)


def test_sync_batchnorm_matches_global_batch():
    spawn(_w_syncbn, 2)


def test_sync_batchnorm_4d_single_process_matches_batchnorm():
    torch.manual_seed(0)
    bn = nn.BatchNorm2d(6)
    sbn = convert_sync_batchnorm(copy.deepcopy(bn))
    x = torch.randn(4, 6, 5, 5, requires_grad=True)
    x2 = x.detach().clone().requires_grad_(True)
    y, y2 = bn(x), sbn(x2)
    torch.testing.assert_close(y, y2, rtol=1e-5, atol=1e-5)
    g = torch.randn_like(y)
    y.backward(g)
    y2.backward(g)
    torch.testing.assert_close(x.grad, x2.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(bn.weight.grad, sbn.weight.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(bn.running_var, sbn.running_var, rtol=1e-5, atol=1e-6)


# ----------------------------------------------------------------------------- sharded scaler
def _w_scaler(rank, world):
    model = make_model()
    opt = OSS(model.parameters(), optim=torch.optim.SGD, lr=0.1)
    sddp = ShardedDataParallel(model, opt)
    scaler = ShardedGradScaler(init_scale=2.0 ** 10, growth_interval=2)
    before = [p.detach().clone() for p in model.parameters()]
    x = torch.randn(4, 12)
    loss = sddp(x).sum()
    scaler.scale(loss).backward()
    # poison one gradient on the rank that owns the first parameter only
    first = next(iter(model.parameters()))
    if opt.param_to_rank[first] == rank:
        first.grad[0, 0] = float("inf")
    scaler.unscale_(opt)
    scaler.step(opt)
    scaler.update()
    # every rank must have skipped the step, and the scale backed off
    for a, b in zip(model.parameters(), before):
        assert torch.equal(a, b)
    assert scaler.get_scale() == 2.0 ** 9
    sddp.zero_grad()
    for _ in range(2):
        scaler.scale(sddp(x).sum()).backward()
        scaler.step(opt)
        scaler.update()
        sddp.zero_grad()
    assert scaler.get_scale() == 2.0 ** 10
    sd = scaler.state_dict()
    assert set(sd) == {"scale", "growth_factor", "backoff_factor", "growth_interval", "_growth_tracker"}


def test_sharded_grad_scaler_skips_on_any_rank_inf():
    spawn(_w_scaler, 2)


def _w_exposed_comm(rank, world):
    import torch.nn as nn
    from pytorch_distributedtraining_b200.parallel.ddp import DistributedDataParallel
    from pytorch_distributedtraining_b200.utils.timing import exposed_comm_ms
    torch.manual_seed(0)
    model = DistributedDataParallel(nn.Sequential(nn.Linear(64, 64), nn.ReLU(), nn.Linear(64, 8)))
    x = torch.randn(16, 64)

    def fwd_bwd():
        model.zero_grad(set_to_none=True)
        model(x).square().mean().backward()

    r = exposed_comm_ms(model, fwd_bwd, iters=3, warmup=2)
    assert set(r) == {"sync_ms", "no_sync_ms", "exposed_comm_ms"}
    assert r["sync_ms"] > 0 and r["no_sync_ms"] > 0 and r["exposed_comm_ms"] >= 0
    # the helper must leave the replicas synchronised: gradients equal on every rank
    g = torch.cat([p.grad.flatten() for p in model.parameters()])
    ref = g.clone()
    torch.distributed.broadcast(ref, src=0)
    assert torch.allclose(g, ref)


def test_exposed_comm_helper_runs_under_gloo():
    from pytorch_distributedtraining_b200.utils.testing import spawn
    spawn(_w_exposed_comm, 2)
