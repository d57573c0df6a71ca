"""Peer-memory collectives.

Part 1 runs every collective kernel with all ranks emulated on ONE GPU (cooperative launch,
``ctx.emulate``): same kernel code, same flag protocol, local pointers instead of peer
mappings (B200_PROFILING.md: never run kernels that wait on each other as separate launches
on one GPU).  Part 2 (needs >= 2 GPUs) runs the real thing over NVLink against NCCL.
"""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda")


@pytest.fixture(scope="module")
def C():
    from pytorch_distributedtraining_b200.ops import collectives, require_library
    require_library()
    return collectives


def _rand(world, n, dtype, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(n, generator=g) * scale).to(dtype).to(cuda) for _ in range(world)]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_barrier(C, world):
    em = C.EmulatedPeers(world, max_blocks=8)
    em.ctx.barrier(blocks=8, rounds=50)
    em.ctx.barrier(blocks=3, rounds=7)
    torch.cuda.synchronize()
    assert not em.timed_out()
    assert all(int(e[0]) == 57 for e in em.epochs) and all(int(e[7]) == 50 for e in em.epochs)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("algo,n", [(0, 4096), (1, 4096), (1, 1 << 20), (0, 128), (1, 8 * 1024 + 64)])
def test_emulated_allreduce(C, world, dtype, algo, n):
    em = C.EmulatedPeers(world, max_blocks=8)
    xs = _rand(world, n, dtype, seed=n + world)
    want = torch.stack([x.float() for x in xs]).sum(0) * (0.5 / world)
    mult = [torch.tensor(0.5, device=cuda) for _ in range(world)]
    flags = [torch.zeros((), device=cuda) for _ in range(world)]
    for it in range(2):   # second pass re-uses flags/epochs: the protocol must be re-entrant
        bufs = [x.clone() for x in xs]
        em.ctx.allreduce(C.table(bufs), 0, n, dtype, pre_mult=1.0 / world, mult_dev=C.table(mult),
                         found_inf=C.table(flags), algo=algo, blocks=8)
        torch.cuda.synchronize()
        assert not em.timed_out()
        tol = {torch.float32: 1e-5, torch.float16: 2e-3, torch.bfloat16: 2e-2}[dtype]
        for b in bufs:
            torch.testing.assert_close(b.float(), want, rtol=tol, atol=tol)
            assert torch.equal(b, bufs[0]), "replicas must be bit-identical"
        assert all(f.item() == 0 for f in flags)
    bufs = [x.clone() for x in xs]
    bufs[world - 1][n // 2] = float("inf")
    em.ctx.allreduce(C.table(bufs), 0, n, dtype, pre_mult=1.0, mult_dev=None, found_inf=C.table(flags), algo=algo, blocks=8)
    torch.cuda.synchronize()
    assert all(f.item() == 1 for f in flags), "every rank must see the non-finite value"


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_emulated_reduce_scatter_and_all_gather(C, world, dtype):
    em = C.EmulatedPeers(world, max_blocks=8)
    n_per = 4096 + 64
    xs = _rand(world, n_per * world, dtype, seed=3)
    outs = [torch.full((n_per,), 1.0, device=cuda) for _ in range(world)]
    em.ctx.reduce_scatter(C.table(xs), 0, n_per, dtype, C.table(outs), pre_mult=1.0 / world, accumulate=True, blocks=8)
    torch.cuda.synchronize()
    total = torch.stack([x.float() for x in xs]).sum(0) / world
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    for r in range(world):
        torch.testing.assert_close(outs[r], 1.0 + total[r * n_per:(r + 1) * n_per], rtol=tol, atol=tol)
    # all-gather with the fp32 -> dtype cast fused in
    shards = _rand(world, n_per, torch.float32, seed=4)
    fulls = [torch.zeros(n_per * world, dtype=dtype, device=cuda) for _ in range(world)]
    em.ctx.all_gather(C.table(fulls), 0, C.table(shards), n_per, torch.float32, dtype, blocks=8)
    torch.cuda.synchronize()
    want = torch.cat(shards).to(dtype)
    for f in fulls:
        assert torch.equal(f, want)
    assert not em.timed_out()


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("wire", [torch.float32, torch.float16])
@pytest.mark.parametrize("prereduced", [False, True])
def test_emulated_zero_adam_matches_torch_adamw(C, world, wire, prereduced):
    from pytorch_distributedtraining_b200.ops import fused
    em = C.EmulatedPeers(world, max_blocks=8)
    sizes = [4096 + 8 * r for r in range(world)]            # uneven owner segments (multiples of 8)
    seg = [0]
    for s in sizes:
        seg.append(seg[-1] + s)
    total = seg[-1]
    g = torch.Generator().manual_seed(11)
    p0 = torch.randn(total, generator=g).to(cuda)
    ref = p0.clone().requires_grad_(True)
    opt = torch.optim.AdamW([ref], lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=1e-2)
    G = [torch.zeros(total, device=cuda) for _ in range(world)]
    P = [p0.clone().to(wire) if wire != torch.float32 else p0.clone() for _ in range(world)]
    master = [p0[seg[r]:seg[r + 1]].clone() for r in range(world)]
    m = [torch.zeros(sizes[r], device=cuda) for r in range(world)]
    v = [torch.zeros(sizes[r], device=cuda) for r in range(world)]
    step = [torch.zeros((), device=cuda) for _ in range(world)]
    hp = fused.AdamHParams(1e-2, 0.9, 0.99, 1e-8, 1e-2, 1, 0, None)
    for it in range(3):
        grads = _rand(world, total, torch.float32, seed=100 + it)
        for r in range(world):
            G[r].copy_(grads[r])
        ref.grad = torch.stack(grads).sum(0) / world
        opt.step()
        for r in range(world):
            fused.step_advance(step[r], None)
        local = None
        if prereduced:
            local = [(torch.stack(grads).sum(0) / world)[seg[r]:seg[r + 1]].contiguous() for r in range(world)]
        em.ctx.zero_adam(None if prereduced else C.table(G), 0, C.table(P), 0, seg, wire, C.table(master), C.table(m),
                         C.table(v), hp, C.table(step), pre_mult=1.0 if prereduced else 1.0 / world,
                         prereduced=prereduced, local_grad=C.table(local) if prereduced else None, blocks=8)
        torch.cuda.synchronize()
    assert not em.timed_out()
    for r in range(world):
        torch.testing.assert_close(master[r], ref.detach()[seg[r]:seg[r + 1]], rtol=2e-5, atol=2e-6)
        if wire == torch.float32:
            torch.testing.assert_close(P[r], ref.detach(), rtol=2e-5, atol=2e-6)
        else:
            torch.testing.assert_close(P[r].float(), ref.detach(), rtol=2e-3, atol=2e-3)
        assert torch.equal(P[r], P[0])


@pytest.mark.parametrize("world", [2, 8])
def test_emulated_reduce_to_owner_and_small_allreduce(C, world):
    em = C.EmulatedPeers(world, max_blocks=8)
    sizes = [2048 + 8 * r for r in range(world)]
    seg = [0]
    for s in sizes:
        seg.append(seg[-1] + s)
    G = _rand(world, seg[-1], torch.float32, seed=5)
    outs = [torch.zeros(sizes[r], device=cuda) for r in range(world)]
    norm = [torch.zeros(2, device=cuda) for _ in range(world)]
    parts = [torch.zeros(8, device=cuda) for _ in range(world)]
    ctr = [torch.zeros(1, dtype=torch.int32, device=cuda) for _ in range(world)]
    em.ctx.reduce_to_owner(C.table(G), seg, C.table(outs), 1.0 / world, None, C.table(norm), C.table(parts),
                           C.table(ctr), blocks=8)
    torch.cuda.synchronize()
    mean = torch.stack(G).sum(0) / world
    for r in range(world):
        torch.testing.assert_close(outs[r], mean[seg[r]:seg[r + 1]], rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(norm[r][0], (mean[seg[r]:seg[r + 1]] ** 2).sum(), rtol=1e-4, atol=1e-5)
        assert norm[r][1].item() == 0 and ctr[r].item() == 0
    # small all-reduce, several calls in a row (slot double-buffering)
    slot = 1024
    staging = [torch.zeros(2 * slot, device=cuda) for _ in range(world)]
    calls = [torch.zeros(1, dtype=torch.int32, device=cuda) for _ in range(world)]
    for k in range(5):
        io = _rand(world, 257, torch.float32, seed=20 + k)
        want = torch.stack(io).sum(0)
        em.ctx.small_allreduce(C.table(staging), C.table(io), 257, slot, C.table(calls))
        torch.cuda.synchronize()
        for x in io:
            torch.testing.assert_close(x, want, rtol=1e-5, atol=1e-5)
    assert not em.timed_out() and all(c.item() == 5 for c in calls)


@pytest.mark.parametrize("world", [2, 8])
def test_emulated_broadcast(C, world):
    em = C.EmulatedPeers(world, max_blocks=8)
    nbytes = 16 * 4099
    bufs = [torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device=cuda) for _ in range(world)]
    want = bufs[1].clone()
    em.ctx.broadcast(C.table(bufs), 0, nbytes, root=1, blocks=4)
    torch.cuda.synchronize()
    assert not em.timed_out()
    for b in bufs:
        assert torch.equal(b, want)


# ----------------------------------------------------------------------------- real multi-GPU
needs2 = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")


def _w_real(rank, world):
    import torch.distributed as dist
    from pytorch_distributedtraining_b200.ops import collectives as C
    from pytorch_distributedtraining_b200.parallel.symm import PeerArena, SmallAllReduce
    dev = torch.device("cuda", rank)
    arena = PeerArena(dist.group.WORLD, dev, 256 << 20, max_blocks=160, timeout_s=20.0)
    ctx = arena.context("main")
    ctx.barrier(blocks=16, rounds=10)
    for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 2e-2), (torch.float16, 2e-3)):
        for n in (1024, 1 << 16, (1 << 22) + 1024):
            for use_mc, algo in ((False, 0), (False, 1), (True, 1)):
                if use_mc and not arena.mc_base:
                    continue
                sb = arena.alloc(n, dtype)
                g = torch.Generator().manual_seed(1000 + rank)
                x = torch.randn(n, generator=g).to(dtype).to(dev)
                sb.tensor.copy_(x)
                ref = x.float().clone()
                dist.all_reduce(ref)
                flag = torch.zeros((), device=dev)
                torch.cuda.synchronize()
                dist.barrier()
                ctx.allreduce(sb.table, sb.mc if use_mc else 0, n, dtype, pre_mult=1.0, found_inf=C.local_table(rank, flag),
                              algo=algo, blocks=32)
                torch.cuda.synchronize()
                torch.testing.assert_close(sb.tensor.float(), ref, rtol=tol, atol=tol * 4)
                assert flag.item() == 0
    # a non-finite value on ONE rank must raise the flag on EVERY rank, also when the slice holding it is reduced
    # in the switch and only its owner sees the result (NVLS)
    n = 1 << 20
    sb = arena.alloc(n, torch.float32)
    for use_mc, algo in ((False, 0), (False, 1), (True, 1)):
        if use_mc and not arena.mc_base:
            continue
        for bad_rank, pos in ((0, n - 5), (world - 1, 3)):
            sb.tensor.fill_(1.0)
            if rank == bad_rank:
                sb.tensor[pos] = float("inf")
            flag = torch.zeros((), device=dev)
            torch.cuda.synchronize(); dist.barrier()
            ctx.allreduce(sb.table, sb.mc if use_mc else 0, n, torch.float32, pre_mult=1.0,
                          found_inf=C.local_table(rank, flag), algo=algo, blocks=32)
            torch.cuda.synchronize()
            assert flag.item() == 1.0, (use_mc, algo, bad_rank, rank)
            # and the next clean call leaves a fresh flag untouched (the notice rows are epoch-tagged)
            sb.tensor.fill_(1.0)
            flag.zero_()
            torch.cuda.synchronize(); dist.barrier()
            ctx.allreduce(sb.table, sb.mc if use_mc else 0, n, torch.float32, pre_mult=1.0,
                          found_inf=C.local_table(rank, flag), algo=algo, blocks=32)
            torch.cuda.synchronize()
            assert flag.item() == 0.0 and torch.all(sb.tensor == world)
    # reduce-scatter / all-gather
    n_per = 1 << 18
    sb = arena.alloc(n_per * world, torch.bfloat16)
    x = torch.randn(n_per * world, device=dev).bfloat16()
    sb.tensor.copy_(x)
    ref = x.float().clone()
    dist.all_reduce(ref)
    out = torch.zeros(n_per, device=dev)
    for mc in ([0, sb.mc] if arena.mc_base else [0]):
        ctx.reduce_scatter(sb.table, mc, n_per, torch.bfloat16, C.local_table(rank, out), pre_mult=1.0 / world)
        torch.cuda.synchronize()
        torch.testing.assert_close(out, ref[rank * n_per:(rank + 1) * n_per] / world, rtol=2e-2, atol=2e-2)
    full = arena.alloc(n_per * world, torch.bfloat16)
    shard = torch.randn(n_per, device=dev)
    gathered = [torch.zeros(n_per, device=dev) for _ in range(world)]
    dist.all_gather(gathered, shard)
    for mc in ([0, full.mc] if arena.mc_base else [0]):
        full.tensor.zero_()
        torch.cuda.synchronize(); dist.barrier()
        ctx.all_gather(full.table, mc, C.local_table(rank, shard), n_per, torch.float32, torch.bfloat16)
        torch.cuda.synchronize()
        assert torch.equal(full.tensor, torch.cat(gathered).bfloat16())
    small = SmallAllReduce(arena)
    for k in range(4):
        v = torch.full((300,), float(rank + k), device=dev)
        assert small(v, None)
        torch.cuda.synchronize()
        assert torch.all(v == sum(r + k for r in range(world)))
    # W4: all-gather of row-sharded weights fused into the GEMM, shards read over NVLink
    from pytorch_distributedtraining_b200.ops.gemm import gemm_gather
    N_, K_, M_ = 128 * 2 * world, 512, 2048
    rpr = N_ // world
    g = torch.Generator().manual_seed(5)
    w_all = (torch.randn(N_, K_, generator=g) * 0.5).bfloat16().to(dev)
    xg = (torch.randn(M_, K_, generator=g) * 0.5).bfloat16().to(dev)
    shard = arena.alloc(rpr * K_, torch.bfloat16)
    for it in range(2):
        shard.tensor.copy_((w_all[rank * rpr:(rank + 1) * rpr] * (it + 1)).reshape(-1))
        torch.cuda.synchronize(); dist.barrier()
        w_full = torch.zeros(N_, K_, dtype=torch.bfloat16, device=dev)
        y = gemm_gather(xg, shard.ptrs, w_full, rank, world, rpr, ctx=ctx)
        torch.cuda.synchronize()
        assert torch.equal(w_full, w_all * (it + 1))
        torch.testing.assert_close(y.float(), xg.float() @ (w_all * (it + 1)).float().t(), rtol=2e-2, atol=0.3)
        dist.barrier()
    assert not arena.timed_out()
    if rank == 0:
        print(f"[real] world={world} multicast={'yes' if arena.mc_base else 'no'} all peer collectives match NCCL")


@needs2
def test_real_peer_collectives_match_nccl():
    from pytorch_distributedtraining_b200.utils.testing import spawn
    spawn(_w_real, min(torch.cuda.device_count(), 8), backend="nccl")


def _w_facade_peer(rank, world, mode):
    """Same training run twice: peer-memory collectives vs NCCL collectives."""
    import torch.nn as nn
    from pytorch_distributedtraining_b200 import (AMPConfig, B200Config, ClipGradNormConfig, DDPConfig,
                                                  FairscaleFSDPConfig, FairscaleOSSConfig, Stoke, StokeOptimizer)
    from pytorch_distributedtraining_b200.models import resnet18, vit_tiny
    from pytorch_distributedtraining_b200.ops import dispatch
    from pytorch_distributedtraining_b200.ops.loss import cross_entropy
    # two separate runs are compared, and AdamW turns 1e-7 differences in near-zero gradients into lr-sized
    # parameter differences: keep the compute path run-to-run reproducible (the conv -> BatchNorm statistics
    # epilogue accumulates with float atomics) so that what is compared is the collectives
    dispatch.disable("convstats")

    wire16 = mode == "ddp_bf16wire"      # DDP bucket all-reduce with a 16-bit wire (DDPConfig.comm_dtype)
    if wire16:
        mode = "ddp"
    sgd = mode == "sddp_sgd"             # the fused ZeRO-2 step with SGD (momentum + Nesterov + L2) instead of AdamW
    if sgd:
        mode = "sddp"
    ring = mode == "fsdp_ring"
    if ring:
        # four blocks over a ring of two gathered-parameter slots: parameters are re-gathered in backward, the next
        # unit's gather is prefetched, gradients leave through the staged reduce-scatter on the side stream
        import os
        os.environ["B200DP_FSDP_RING"] = "2"
        mode = "fsdp"

    def run(peer):
        torch.manual_seed(0)
        if mode == "fsdp":
            from pytorch_distributedtraining_b200.models.vit import VisionTransformer
            model = VisionTransformer(32, 8, 4, 2, 64, 128, 10) if ring else vit_tiny()
        else:
            model = resnet18(num_classes=10, small_input=True)
        kw = dict(fairscale_oss=mode == "sddp", fairscale_sddp=mode == "sddp", fairscale_fsdp=mode == "fsdp")
        s = Stoke(model=model, verbose=False,
                  optimizer=(StokeOptimizer(optimizer=torch.optim.SGD, optimizer_kwargs={"lr": 0.05, "momentum": 0.9, "nesterov": True,
                                                                                         "weight_decay": 1e-4}) if sgd else
                             StokeOptimizer(optimizer=torch.optim.AdamW, optimizer_kwargs={"lr": 1e-3, "betas": (0.9, 0.99),
                                                                                           "eps": 1e-8, "weight_decay": 1e-4})),
                  loss=cross_entropy, batch_size_per_device=16, gpu=True, fp16="bf16" if mode == "fsdp" else "amp",
                  distributed="ddp", grad_accum_steps=2 if mode == "sddp" else 1,
                  grad_clip=ClipGradNormConfig(max_norm=1.0, norm_type=2.0) if mode != "ddp" else None,
                  configs=[AMPConfig(init_scale=2.0 ** 10),
                           DDPConfig(local_rank=rank, comm_dtype=torch.bfloat16 if (wire16 and peer) else None),
                           FairscaleFSDPConfig(min_num_params=10 ** 9), B200Config(peer_collectives=peer)], **kw)
        assert (s._peer is not None) == peer
        if peer and wire16:
            assert s._peer["ddp"]._wire, "no 16-bit wire buffers were created"
        if peer and mode == "sddp":
            assert "zero" in s._peer, "the fused ZeRO-2 engine did not attach"
        if peer and ring:
            assert not s._peer["fsdp"].resident and len(s._peer["fsdp"].slots) == 2
        s.model_access.train()
        g = torch.Generator().manual_seed(77 + rank)
        losses = []
        for i in range(6):
            x = torch.randn(16, 3, 32, 32, generator=g).cuda()
            y = torch.randint(0, 10, (16,), generator=g).cuda()
            loss = s.loss(s.model(x), y)
            s.backward(loss)
            s.step()
            losses.append(s.detach_and_sync_loss(loss))
        sd = s.wrapped_model.state_dict()      # through the engine: flushes the lazily synchronised BatchNorm statistics
        return losses, {k: v.detach().float().clone() for k, v in sd.items()}

    l_peer, sd_peer = run(True)
    l_nccl, sd_nccl = run(False)
    if rank == 0:
        print(f"[{mode}] peer losses {l_peer}\n[{mode}] nccl losses {l_nccl}")
    for a, b in zip(l_peer, l_nccl):
        assert abs(a - b) < 5e-2 * max(1.0, abs(b)), (l_peer, l_nccl)
    num = sum(((sd_peer[k] - sd_nccl[k]) ** 2).sum() for k in sd_peer if sd_peer[k].is_floating_point())
    den = sum((sd_nccl[k] ** 2).sum() for k in sd_nccl if sd_nccl[k].is_floating_point())
    assert (num / den).sqrt().item() < 2e-2, (num, den)


@needs2
@pytest.mark.parametrize("mode", ["ddp", "ddp_bf16wire", "sddp", "sddp_sgd", "fsdp", "fsdp_ring"])
def test_real_facade_peer_matches_nccl(mode):
    from pytorch_distributedtraining_b200.utils.testing import spawn
    spawn(_w_facade_peer, 2, (mode,), backend="nccl")


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs >= 4 GPUs")
@pytest.mark.parametrize("mode", ["ddp", "sddp", "fsdp_ring"])
def test_real_facade_peer_matches_nccl_on_every_gpu(mode):
    """The same equivalence on all GPUs of the box (4 or 8 ranks): loss trajectories and final weights of the
    peer-memory engines against the NCCL engines."""
    from pytorch_distributedtraining_b200.utils.testing import spawn
    spawn(_w_facade_peer, min(torch.cuda.device_count(), 8), (mode,), backend="nccl")


@pytest.mark.parametrize("world", [2, 8])
def test_emulated_collectives_under_random_rank_skew(C, world):
    """Flag-protocol stress (sanitizers do not see cross-rank races): with ctx.skew_ns set, every (rank, CTA)
    idles a pseudo-random time before each barrier, so ranks arrive in a different order every round; results
    must stay exact and no barrier may time out."""
    em = C.EmulatedPeers(world, max_blocks=8)
    em.ctx.c.skew_ns = 30_000
    n = 8 * 1024 + 64
    for it in range(6):
        xs = _rand(world, n, torch.float32, seed=50 + it)
        want = torch.stack(xs).sum(0) / world
        for algo in (0, 1):
            bufs = [x.clone() for x in xs]
            em.ctx.allreduce(C.table(bufs), 0, n, torch.float32, pre_mult=1.0 / world, mult_dev=None, found_inf=None,
                             algo=algo, blocks=8)
            torch.cuda.synchronize()
            for b in bufs:
                torch.testing.assert_close(b, want, rtol=1e-5, atol=1e-6)
                assert torch.equal(b, bufs[0])
        n_per = 2048
        ins = _rand(world, n_per * world, torch.bfloat16, seed=70 + it)
        outs = [torch.zeros(n_per, device=cuda) for _ in range(world)]
        em.ctx.reduce_scatter(C.table(ins), 0, n_per, torch.bfloat16, C.table(outs), pre_mult=1.0, accumulate=False, blocks=8)
        torch.cuda.synchronize()
        total = torch.stack([x.float() for x in ins]).sum(0)
        for r in range(world):
            torch.testing.assert_close(outs[r], total[r * n_per:(r + 1) * n_per], rtol=2e-2, atol=2e-2)
    em.ctx.barrier(blocks=8, rounds=200)
    torch.cuda.synchronize()
    assert not em.timed_out()
