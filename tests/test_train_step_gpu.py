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
