"""``Stoke.train_step`` with CUDA-graph capture must train exactly like the eager four-call API."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(graph: bool, accum: int, clip, steps: int = 10, sched: bool = True):
    from pytorch_distributedtraining_b200 import AMPConfig, B200Config, ClipGradNormConfig, Stoke, StokeOptimizer
    from pytorch_distributedtraining_b200.models import resnet18
    from pytorch_distributedtraining_b200.ops.loss import cross_entropy
    torch.manual_seed(0)
    s = Stoke(model=resnet18(num_classes=10, small_input=True), verbose=False,
              optimizer=StokeOptimizer(optimizer=torch.optim.AdamW,
                                       optimizer_kwargs={"lr": 1e-3, "betas": (0.9, 0.99), "eps": 1e-8, "weight_decay": 1e-4}),
              loss=cross_entropy, batch_size_per_device=32, gpu=True, fp16="amp", grad_accum_steps=accum,
              grad_clip=None if clip is None else ClipGradNormConfig(max_norm=clip, norm_type=2.0),
              configs=[AMPConfig(init_scale=2.0 ** 8), B200Config(cuda_graph=graph)])
    s.model_access.train()
    lr_sched = torch.optim.lr_scheduler.StepLR(s.optimizer, step_size=2, gamma=0.5) if sched else None
    g = torch.Generator().manual_seed(5)
    losses = []
    for i in range(steps):
        x = torch.randn(32, 3, 32, 32, generator=g).cuda().contiguous(memory_format=torch.channels_last)
        y = torch.randint(0, 10, (32,), generator=g).cuda()
        if graph:
            loss = s.train_step(x, y)
        else:
            l = s.loss(s.model(x), y)
            s.backward(l)
            s.step()
            loss = s.step_loss
        losses.append(float(loss))
        if lr_sched is not None and (i + 1) % accum == 0:
            lr_sched.step()
    assert s.backward_steps == steps and s.optimizer_steps == steps // accum
    return losses, [p.detach().clone() for p in s.model_access.parameters()], s


@pytest.mark.parametrize("accum,clip", [(1, None), (2, 1.0)])
def test_graphed_train_step_matches_eager(accum, clip):
    l_e, p_e, _ = _run(False, accum, clip)
    l_g, p_g, s = _run(True, accum, clip)
    assert any(e["graph"] is not None for e in s._graphs.values()), "no CUDA graph was captured"
    for a, b in zip(l_e, l_g):
        assert abs(a - b) <= 2e-2 * max(1.0, abs(a)), (l_e, l_g)
    num = sum(((a - b) ** 2).sum() for a, b in zip(p_g, p_e))
    den = sum((b ** 2).sum() for b in p_e)
    # split-K weight gradients are combined with TMA reduce-add (order not fixed) and AdamW turns the
    # rounding noise of near-zero gradients into +-lr moves, so two runs agree to ~1e-2, not bitwise
    assert (num / den).sqrt().item() < 2e-2, (num, den)
    assert float(s.ema_loss) == float(s.ema_loss) and s.ema_loss is not None


def test_optimizer_keeps_16bit_weight_shadows_the_layers_read():
    """The fused AdamW kernel writes the compute-dtype copy of every matrix-shaped parameter; convolutions and linears
    use it instead of casting fp32 weights each forward.  The shadow must equal ``p.to(dtype)`` after every step
    (including a step skipped on overflow), and a parameter modified behind the optimizer's back must NOT use it."""
    from pytorch_distributedtraining_b200 import AMPConfig, Stoke, StokeOptimizer
    from pytorch_distributedtraining_b200.models import resnet18
    from pytorch_distributedtraining_b200.ops import fused
    from pytorch_distributedtraining_b200.ops.loss import cross_entropy
    torch.manual_seed(0)
    s = Stoke(model=resnet18(num_classes=10, small_input=True), verbose=False,
              optimizer=StokeOptimizer(optimizer=torch.optim.AdamW, optimizer_kwargs={"lr": 1e-3}), loss=cross_entropy,
              batch_size_per_device=8, gpu=True, fp16="amp", configs=[AMPConfig(init_scale=2.0 ** 10)])
    s.model_access.train()
    mats = [p for p in s.model_access.parameters() if p.dim() >= 2]
    assert all(getattr(p, fused.LP_ATTR, None) is not None for p in mats)
    x, y = torch.randn(8, 3, 32, 32, device="cuda"), torch.randint(0, 10, (8,), device="cuda")
    for it in range(4):
        loss = s.loss(s.model(x), y)
        s.backward(loss)
        if it == 2:
            mats[0].grad[(0,) * mats[0].dim()] = float("inf")        # this step is skipped on the device
        s.step()
        for p in mats:
            assert torch.equal(getattr(p, fused.LP_ATTR), p.detach().half())
            assert fused.lp_weight(p, torch.float16) is getattr(p, fused.LP_ATTR)
    with torch.no_grad():
        mats[0].mul_(2.0)                                    # e.g. a checkpoint load
    assert fused.lp_weight(mats[0], torch.float16) is not getattr(mats[0], fused.LP_ATTR)
    loss = s.loss(s.model(x), y); s.backward(loss); s.step()
    assert fused.lp_weight(mats[0], torch.float16) is getattr(mats[0], fused.LP_ATTR)


@pytest.mark.parametrize("fp16", ["amp", "bf16"])
def test_oss_clip_through_the_fused_norm_matches_the_per_parameter_loop(fp16):
    """A bare OSS (one GPU, or NCCL without the peer engine) reaches the fused optimizers' one-read norm / inf check
    (``OSS.fused_unscale_clip``).  (1) On the SAME gradients its norm and multiplier equal the fp32 torch computation;
    (2) the training trajectory matches the per-parameter clip loop it replaces (``B200DP_DISABLE=oss_fused_clip``) with
    gradient accumulation and clipping that really clips, fp16 (GradScaler) and bf16.  Adam turns rounding-level
    gradient differences into sign flips of near-zero elements, so (2) is a loss-trajectory check with a loose bound on
    the parameters; (1) is the exact one."""
    import argparse
    import os, sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from pytorch_distributedtraining_b200.ops import dispatch
    from pytorch_distributedtraining_b200.parallel.oss import OSS

    def build():
        torch.manual_seed(0)
        cfg = dict(model="resnet18", batch=32, shape=(3, 32, 32), classes=10, fp16=fp16, oss=True, sddp=True, fsdp=False,
                   accum=2, clip=0.05)
        args = argparse.Namespace(cpu=False, batch_size=0, impl="ours", no_graph=True, sync_bn=False)
        return bench.build_ours(cfg, args, 1)

    def batch(g, device):
        x = torch.randn(32, 3, 32, 32, generator=g).to(device).contiguous(memory_format=torch.channels_last)
        return x, torch.randint(0, 10, (32,), generator=g).to(device)

    def run(fused: bool):
        (dispatch.enable if fused else dispatch.disable)("oss_fused_clip")
        try:
            stoke, step, device = build()
            assert isinstance(stoke.optimizer, OSS) and ("fused_unscale_clip" in stoke.optimizer.__dict__)
            g = torch.Generator(device="cpu").manual_seed(1)
            losses = [float(step(*batch(g, device))) for _ in range(8)]
            params = [p.detach().float().clone() for p in stoke.model_access.parameters()]
        finally:
            dispatch.enable("oss_fused_clip")
        return losses, params

    # (1) the norm itself, on gradients left in place by two accumulated micro-batches (bf16: no loss scale involved)
    if fp16 == "bf16":
        stoke, _step, device = build()
        g = torch.Generator(device="cpu").manual_seed(2)
        for _ in range(2):
            x, y = batch(g, device)
            stoke.backward(stoke.loss(stoke.model(x), y))
        grads = [p.grad for p in stoke.model_access.parameters() if p.grad is not None]
        want = torch.sqrt(sum((gr.float() ** 2).sum() for gr in grads))
        before = [gr.clone() for gr in grads]
        total = stoke.optimizer.clip_grad_norm(0.05, 2.0)
        torch.testing.assert_close(total.float().reshape(()), want.reshape(()), rtol=1e-3, atol=1e-6)
        inner = stoke.optimizer.optim
        torch.testing.assert_close(inner._grad_mult.reshape(()), (0.05 / (want + 1e-6)).clamp(max=1.0).reshape(()),
                                   rtol=1e-3, atol=1e-7)
        assert all(torch.equal(a, b) for a, b in zip(before, grads))      # gradients are not rewritten
        stoke.optimizer.step()                                            # consumes the multiplier without error
        assert all(torch.isfinite(p).all() for p in stoke.model_access.parameters())

    # (2) trajectories
    l_new, p_new = run(True)
    l_old, p_old = run(False)
    assert all(l == l for l in l_new)
    torch.testing.assert_close(torch.tensor(l_new), torch.tensor(l_old), rtol=3e-2, atol=3e-2)
    num = sum(((a - b) ** 2).sum() for a, b in zip(p_new, p_old))
    den = sum((b ** 2).sum() for b in p_old)
    assert float((num / den).sqrt()) < 8e-2
