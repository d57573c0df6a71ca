"""Facade semantics on CPU: flag validation, accumulation gating, clip dispatch, loss
bookkeeping, DataLoader contract, checkpoint layout (SURVEY §4 unit + checkpoint tiers)."""
import os

import pytest
import torch
import torch.nn as nn

from pytorch_distributedtraining_b200 import (AMPConfig, ClipGradConfig, ClipGradNormConfig, DDPConfig,
                                              DistributedOptions, FairscaleOSSConfig, FP16Options, Stoke,
                                              StokeOptimizer, SyntheticImageDataset)
from pytorch_distributedtraining_b200.data import DistributedSampler
from pytorch_distributedtraining_b200.utils.testing import spawn

ADAMW = {"lr": 1e-2, "betas": (0.9, 0.99), "eps": 1e-8, "weight_decay": 1e-4}


def tiny():
    torch.manual_seed(3)
    return nn.Sequential(nn.Linear(8, 16), nn.ReLU(), nn.Linear(16, 4))


def build(**kw):
    args = dict(model=tiny(), verbose=False, optimizer=StokeOptimizer(optimizer=torch.optim.AdamW, optimizer_kwargs=ADAMW),
                loss=nn.MSELoss(), batch_size_per_device=4)
    args.update(kw)
    return Stoke(**args)


def test_flag_validation():
    with pytest.raises(ValueError):
        build(fairscale_sddp=True, distributed="ddp")           # sddp needs oss
    with pytest.raises(ValueError):
        build(fairscale_oss=True)                                # fairscale needs ddp
    with pytest.raises(ValueError):
        build(fp16="amp")                                        # amp needs gpu
    with pytest.raises(ValueError):
        build(fairscale_fsdp=True, fairscale_oss=True, distributed="ddp")
    with pytest.raises(NotImplementedError):
        build(distributed=DistributedOptions.horovod.value)
    with pytest.raises(NotImplementedError):
        build(fp16=FP16Options.apex_O1)
    with pytest.raises(ValueError):
        build(distributed="mpi")
    with pytest.raises(TypeError):
        build(configs=[object()])
    with pytest.raises(ValueError):
        build(configs=[AMPConfig(), AMPConfig()])
    assert DistributedOptions.ddp.value == "ddp" and FP16Options.amp.value == "amp"


def test_step_is_noop_until_accumulation_boundary():
    s = build(grad_accum_steps=2, grad_clip=ClipGradNormConfig(max_norm=0.1, norm_type=2.0))
    s.model_access.train()
    x, y = torch.randn(4, 8), torch.randn(4, 4)
    before = [p.detach().clone() for p in s.model_access.parameters()]
    loss = s.loss(s.model(x), y)
    raw = nn.functional.mse_loss(s.model_access(x), y)
    torch.testing.assert_close(loss, raw / 2)                   # training: divided by accum steps
    s.backward(loss)
    s.step()                                                     # micro-step 1: no-op
    assert s.optimizer_steps == 0 and s.backward_steps == 1
    for a, b in zip(s.model_access.parameters(), before):
        assert torch.equal(a, b)
    s.backward(s.loss(s.model(x), y))
    s.step()                                                     # boundary
    assert s.optimizer_steps == 1 and s.backward_steps == 2
    assert any(not torch.equal(a, b) for a, b in zip(s.model_access.parameters(), before))
    assert all(p.grad is None for p in s.model_access.parameters())
    # eval mode: raw loss, counters and EMA untouched
    ema = s.ema_loss.clone()
    s.model_access.eval()
    with torch.no_grad():
        v = s.loss(s.model(x), y)
    torch.testing.assert_close(v, nn.functional.mse_loss(s.model_access(x), y))
    assert torch.equal(s.ema_loss, ema)
    assert isinstance(s.detach_and_sync_loss(v), float)
    assert s.effective_batch_size == 8 and s.grad_accum == 2 and s.batch_size == 4
    assert s.rank == 0 and s.world_size == 1


def test_accumulated_step_equals_big_batch_step():
    s = build(grad_accum_steps=2)
    ref = tiny()
    opt = torch.optim.AdamW(ref.parameters(), **ADAMW)
    s.model_access.train()
    x, y = torch.randn(8, 8), torch.randn(8, 4)
    for i in range(2):
        s.backward(s.loss(s.model(x[i * 4:(i + 1) * 4]), y[i * 4:(i + 1) * 4]))
        s.step()
    nn.functional.mse_loss(ref(x), y).backward()
    opt.step()
    for a, b in zip(s.model_access.parameters(), ref.parameters()):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)


def test_value_clip_and_ema():
    s = build(grad_clip=ClipGradConfig(clip_value=1e-9), ema_weight=0.5)
    s.model_access.train()
    x, y = torch.randn(4, 8), torch.randn(4, 4)
    l1 = s.loss(s.model(x), y)
    s.backward(l1)
    s.step()
    l2 = s.loss(s.model(x), y)
    torch.testing.assert_close(s.ema_loss, 0.5 * l1.detach() + 0.5 * l2.detach())
    s.print_ema_loss("EMA")
    s.print_on_devices("hello")
    s.reset()
    s.reset_ema()
    assert s.ema_loss is None


def test_dataloader_contract():
    s = build()
    ds = SyntheticImageDataset(20, shape=(3, 8, 8), num_classes=5)
    with pytest.raises(ValueError):
        s.DataLoader(ds, batch_size=3)
    dl = s.DataLoader(dataset=ds, num_workers=0, multiprocessing_context="spawn")
    xb, yb = next(iter(dl))
    assert xb.shape == (4, 3, 8, 8) and yb.shape == (4,)
    sm = DistributedSampler(ds, num_replicas=4, rank=1, shuffle=False)
    assert list(sm) == list(range(1, 20, 4))
    a = list(DistributedSampler(ds, num_replicas=2, rank=0, seed=1))
    smp = DistributedSampler(ds, num_replicas=2, rank=0, seed=1)
    e0, e1 = list(smp), list(smp)
    assert e0 == a and e0 != e1, "auto epoch advance reshuffles between epochs"


def test_checkpoint_layout_and_resume(tmp_path):
    s = build(grad_accum_steps=1)
    s.model_access.train()
    x, y = torch.randn(4, 8), torch.randn(4, 4)
    for _ in range(3):
        s.backward(s.loss(s.model(x), y))
        s.step()
    path, tag = s.save(path=str(tmp_path), name="model_0_0.10_0.20", extras={"epoch": 7})
    assert tag == "stoke-model_0_0.10_0.20-backward-step-3"
    ck = torch.load(os.path.join(path, tag + ".pt"), weights_only=False)
    assert {"backward_step", "grad_accum_step", "optimizer_step", "stoke_status", "model_state_dict",
            "optimizer_state_dict", "scaler_state_dict", "extras"} <= set(ck)
    plain = tiny()
    plain.load_state_dict(ck["model_state_dict"], strict=True)
    torch.optim.AdamW(plain.parameters(), **ADAMW).load_state_dict(ck["optimizer_state_dict"])
    s2 = build()
    extras = s2.load(str(tmp_path), tag)
    assert extras == {"epoch": 7} and s2.backward_steps == 3 and s2.optimizer_steps == 3
    # the resumed object continues identically
    for obj in (s, s2):
        obj.model_access.train()
        obj.backward(obj.loss(obj.model(x), y))
        obj.step()
    for a, b in zip(s.model_access.parameters(), s2.model_access.parameters()):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)


# ----------------------------------------------------------------------------- distributed facade
def _w_facade(rank, world, mode, tmp):
    kw = dict(distributed=DistributedOptions.ddp.value, grad_accum_steps=2,
              grad_clip=ClipGradNormConfig(max_norm=0.1, norm_type=2.0),
              configs=[AMPConfig(init_scale=2.0 ** 14), DDPConfig(local_rank=rank, convert_to_sync_batch_norm=True,
                                                                    backend="gloo"),
                       FairscaleOSSConfig(broadcast_fp16=False)])
    if mode == "sddp":
        kw.update(fairscale_oss=True, fairscale_sddp=True)
    elif mode == "oss":
        kw.update(fairscale_oss=True)
    elif mode == "fsdp":
        kw.update(fairscale_fsdp=True)
    s = build(**kw)
    assert s.world_size == world and s.rank == rank
    ds = SyntheticImageDataset(32, shape=(8,), num_classes=4)
    g = torch.Generator().manual_seed(5)
    X, Y = torch.randn(4, 8, 8, generator=g), torch.randn(4, 8, 4, generator=g)
    s.model_access.train()
    for i in range(4):
        xb, yb = X[i, rank * 4:(rank + 1) * 4], Y[i, rank * 4:(rank + 1) * 4]
        loss = s.loss(s.model(xb), yb)
        s.backward(loss)
        s.step()
        synced = s.detach_and_sync_loss(loss)
        assert isinstance(synced, float)
    assert s.optimizer_steps == 2 and s.backward_steps == 4
    # oracle: single-process, global batch, accumulate 2, clip 0.1
    ref = tiny()
    opt = torch.optim.AdamW(ref.parameters(), **ADAMW)
    for i in range(4):
        (nn.functional.mse_loss(ref(X[i]), Y[i]) / 2).backward()
        if i % 2 == 1:
            torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.1)
            opt.step()
            opt.zero_grad()
    path, tag = s.save(path=tmp, name=f"m_{mode}")
    ck = torch.load(os.path.join(path, tag + ".pt"), weights_only=False)
    for (k, v), p in zip(ck["model_state_dict"].items(), ref.parameters()):
        torch.testing.assert_close(v, p.detach(), rtol=1e-5, atol=1e-6)
    if mode != "fsdp":
        plain = torch.optim.AdamW(tiny().parameters(), **ADAMW)
        plain.load_state_dict(ck["optimizer_state_dict"])
        ref_sd = opt.state_dict()
        for k in ref_sd["state"]:
            torch.testing.assert_close(ck["optimizer_state_dict"]["state"][k]["exp_avg"], ref_sd["state"][k]["exp_avg"],
                                       rtol=1e-4, atol=1e-6)
    s2 = build(**kw)
    s2.load(tmp, tag)
    assert s2.optimizer_steps == 2
    for obj in (s, s2):
        obj.model_access.train()
        for i in range(2):
            obj.backward(obj.loss(obj.model(X[i, rank * 4:(rank + 1) * 4]), Y[i, rank * 4:(rank + 1) * 4]))
            obj.step()
    sd1 = s.wrapped_model.state_dict() if mode == "fsdp" else s.model_access.state_dict()
    sd2 = s2.wrapped_model.state_dict() if mode == "fsdp" else s2.model_access.state_dict()
    for k in sd1:
        torch.testing.assert_close(sd1[k], sd2[k], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("mode", ["ddp", "oss", "sddp", "fsdp"])
def test_facade_distributed_matches_single_process(mode, tmp_path):
    spawn(_w_facade, 2, (mode, str(tmp_path)))
