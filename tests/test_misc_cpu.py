"""Data path, launcher, logging, metrics, compat shims and the example scripts (CPU)."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_metrics():
    from pytorch_distributedtraining_b200.utils import metrics
    a, b = torch.zeros(2, 3, 4, 4), torch.full((2, 3, 4, 4), 0.1)
    assert abs(float(metrics.mae(a, b)) - 0.1) < 1e-6
    assert abs(float(metrics.psnr(a, b)) - 20.0) < 1e-3
    assert float(metrics.psnr(b, b)) > 90
    logits = torch.tensor([[0.1, 0.9], [0.8, 0.2]])
    assert float(metrics.accuracy(logits, torch.tensor([1, 1]))) == 0.5
    assert float(metrics.reduce_mean(torch.tensor(3.0))) == 3.0


def test_paired_image_dataset(tmp_path):
    from PIL import Image
    from pytorch_distributedtraining_b200 import CustomDataset
    lr, hr = tmp_path / "lr", tmp_path / "hr"
    lr.mkdir(); hr.mkdir()
    for i in range(3):
        Image.new("RGB", (8, 8), (i * 40, 0, 0)).save(lr / f"{i}.png")
        Image.new("RGB", (16, 16), (i * 40, 0, 0)).save(hr / f"{i}.png")
    ds = CustomDataset(str(lr), str(hr))
    x, y = ds[2]
    assert len(ds) == 3 and x.shape == (3, 8, 8) and y.shape == (3, 16, 16)
    assert abs(float(x[0, 0, 0]) - 80 / 255) < 1e-6 and float(x.max()) <= 1.0


def test_bucketed_sampler_partitions_without_overlap():
    from pytorch_distributedtraining_b200 import BucketedDistributedSampler
    ds = list(range(64))
    seen = []
    for r in range(2):
        s = BucketedDistributedSampler(ds, buckets=2, batch_size=4, sorted_idx=list(range(64)), num_replicas=2, rank=r,
                                       shuffle=True, seed=3)
        idx = list(s)
        assert len(idx) == len(s) == 32
        seen += idx
    assert sorted(seen) == list(range(64))


def test_run_logger_offline(tmp_path, monkeypatch):
    from pytorch_distributedtraining_b200.utils.logging import RunLogger, train_log, val_log
    monkeypatch.setenv("WANDB_MODE", "disabled")
    lg = RunLogger(project="t", config={"a": 1}, log_dir=str(tmp_path))
    train_log(lg, torch.tensor(0.5), 10, 0)
    val_log(lg, 0.4, 0.1, 30.0, 10, 0)
    lg.finish()
    lines = [json.loads(l) for l in open(lg.path)]
    assert lines[0]["event"] == "init" and lines[1]["train_loss"] == 0.5 and lines[2]["PSNR"] == 30.0
    assert RunLogger(project="t", rank=1, log_dir=str(tmp_path)).log({"x": 1}) is None   # non-zero ranks stay silent


def test_env_rank_info_tolerates_the_reference_none_string(monkeypatch):
    from pytorch_distributedtraining_b200.launch import env_rank_info, find_free_port
    monkeypatch.setenv("LOCAL_RANK", "None")           # what Stoke-DDP.py:153 writes when the variable is unset
    monkeypatch.delenv("RANK", raising=False)
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    assert env_rank_info() == (0, 0, 1)
    assert 1024 < int(find_free_port()) < 65536


def test_compat_shims_resolve_reference_imports():
    code = ("import sys; sys.path.insert(0, 'compat'); "
            "from stoke import DeepspeedConfig, DeepspeedZeROConfig, Stoke, StokeOptimizer, AMPConfig, ClipGradNormConfig, "
            "DDPConfig, DistributedOptions, FairscaleOSSConfig, FP16Options; "
            "from fairscale.optim.oss import OSS; from fairscale.nn.data_parallel import ShardedDataParallel as ShardedDDP; "
            "print(Stoke.__name__, OSS.__name__, ShardedDDP.__name__)")
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert out.stdout.split() == ["Stoke", "OSS", "ShardedDataParallel"]


def test_kernel_library_builds_and_exports_the_header_api():
    """Every function declared in include/b200dp.h is exported by the built library."""
    import re
    from pytorch_distributedtraining_b200.ops import build_library
    so = build_library()
    header = open(os.path.join(ROOT, "include", "b200dp.h")).read()
    declared = set(re.findall(r"\b(b200dp_[a-z0-9_]+)\s*\(", header)) - {"b200dp_dtype"}
    syms = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(b200dp_[a-z0-9_]+)\b", syms))
    assert declared and declared <= exported, sorted(declared - exported)


@pytest.mark.parametrize("script,args", [
    ("fairscale_ddp.py", ["--world-size", "2", "--epochs", "1", "--samples", "40", "--batch-size", "4"]),
    ("stoke_ddp.py", ["--nEpochs", "1", "--samples", "12", "--batchSize", "2"]),
])
def test_example_scripts_run_offline(script, args, tmp_path):
    env = dict(os.environ, WANDB_MODE="disabled")
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", script)] + args, cwd=str(tmp_path), env=env,
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-3000:]
    if script == "stoke_ddp.py":
        assert os.path.isdir(tmp_path / "checkpoint") and any(f.endswith(".pt") for f in os.listdir(tmp_path / "checkpoint"))


def test_conv_bn_pairing_and_carry_fallbacks_on_cpu():
    """The kernel-path hooks must be inert off the GPU: pairing only sets a flag, forward_carry returns the
    input itself, and a bottleneck block gives the same gradients as the textbook formulation."""
    import torch.nn as nn
    import torch.nn.functional as F
    from pytorch_distributedtraining_b200.models.resnet import Bottleneck
    from pytorch_distributedtraining_b200.ops.layers import BatchNormAct2d, Conv2d, LayerNorm, Linear, pair_conv_bn
    torch.manual_seed(0)
    blk = Bottleneck(16, 4)
    assert pair_conv_bn(blk) == 3 and blk.conv1.emit_bn_stats and blk.conv3.emit_bn_stats
    x = torch.randn(2, 16, 8, 8, requires_grad=True)
    y = blk(x)
    y.square().mean().backward()
    gx = x.grad.clone()
    x.grad = None
    out = blk.bn1(blk.conv1(x))
    out = blk.bn2(blk.conv2(out))
    ref = blk.bn3(blk.conv3(out), residual=x)
    torch.testing.assert_close(y, ref)
    ref.square().mean().backward()
    torch.testing.assert_close(gx, x.grad)
    c = Conv2d(4, 4, 1)
    t = torch.randn(1, 4, 3, 3)
    yc, same = c.forward_carry(t)
    assert same is t and torch.equal(yc, c(t))
    ln = LayerNorm(8)
    u = torch.randn(3, 8)
    yl, same = ln.forward_carry(u)
    assert same is u and torch.equal(yl, ln(u))
    lin = Linear(8, 8)
    torch.testing.assert_close(lin(u, residual=u), F.linear(u, lin.weight, lin.bias) + u)


def test_attention_layout_and_host_layout_helpers_on_cpu():
    import torch.nn.functional as F
    from pytorch_distributedtraining_b200.data import pin_channels_last
    from pytorch_distributedtraining_b200.ops.attention import merge_heads, split_qkv
    torch.manual_seed(1)
    B, S, H, d = 2, 5, 3, 8
    x = torch.randn(B, S, 3 * H * d, requires_grad=True)
    q, k, v = split_qkv(x, H)
    y = merge_heads(F.scaled_dot_product_attention(q, k, v))
    y.square().sum().backward()
    g1, x.grad = x.grad.clone(), None
    qkv = x.view(B, S, 3, H, d).permute(2, 0, 3, 1, 4)
    y2 = F.scaled_dot_product_attention(qkv[0], qkv[1], qkv[2]).transpose(1, 2).reshape(B, S, H * d)
    y2.square().sum().backward()
    torch.testing.assert_close(y, y2)
    torch.testing.assert_close(g1, x.grad)
    img = torch.randn(2, 3, 4, 5)
    p = pin_channels_last(img)
    assert torch.equal(p, img) and p.is_contiguous(memory_format=torch.channels_last)


def test_dataloader_collates_channels_last_on_the_host():
    from pytorch_distributedtraining_b200.data import StokeDataLoader, SyntheticImageDataset
    ds = SyntheticImageDataset(16, shape=(3, 8, 8), num_classes=4)
    plain = next(iter(StokeDataLoader(ds, batch_size=4)))
    cl = next(iter(StokeDataLoader(ds, batch_size=4, channels_last=True)))
    assert torch.equal(plain[0], cl[0]) and torch.equal(plain[1], cl[1])
    assert cl[0].is_contiguous(memory_format=torch.channels_last) and not plain[0].is_contiguous(memory_format=torch.channels_last)


def test_c_host_links_against_the_kernel_library(tmp_path):
    """The kernel library is a plain C ABI: a C program compiles against include/b200dp.h, links the in-tree
    .so and can plan (scratch sizes, envelopes) and get error text without Python, torch or a GPU."""
    import os, shutil, subprocess
    from pytorch_distributedtraining_b200.ops import _lib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    so = _lib.lib_path()
    if not os.path.exists(so):
        _lib.build_library()
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None or not os.path.isdir("/usr/local/cuda/lib64"):
        import pytest
        pytest.skip("no C compiler / CUDA runtime library on this box")
    exe = str(tmp_path / "b200dp_host")
    cmd = [cc, "-I", os.path.join(root, "include"), os.path.join(root, "examples", "c_host", "main.c"),
           "-L", os.path.dirname(so), "-lb200dp", "-L", "/usr/local/cuda/lib64", "-lcudart",
           "-Wl,-rpath," + os.path.dirname(so), "-Wl,-rpath,/usr/local/cuda/lib64", "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "b200dp version" in r.stdout and "bad arguments" in r.stdout


def test_step_watchdog_fires_only_when_steps_stop():
    import io, time
    from pytorch_distributedtraining_b200.utils.watchdog import StepWatchdog
    hits, log = [], io.StringIO()
    with StepWatchdog(0.4, on_timeout=hits.append, stream=log) as wd:
        for _ in range(8):                      # steady progress: never fires
            time.sleep(0.1)
            wd.beat()
        assert wd.fired == 0 and wd.beats == 8
        wd.pause()
        time.sleep(0.7)                         # paused (e.g. checkpointing): silent
        assert wd.fired == 0
        wd.resume()
        time.sleep(0.9)                         # stuck: fires, dumps stacks, calls back
        assert wd.fired >= 1 and hits and hits[0] >= 0.4
    assert "no training step finished" in log.getvalue() and "Thread" in log.getvalue()


def test_facade_starts_the_step_watchdog_on_request():
    import torch.nn as nn
    from pytorch_distributedtraining_b200 import B200Config, Stoke, StokeOptimizer
    model = nn.Linear(4, 2)
    s = Stoke(model=model, verbose=False, optimizer=StokeOptimizer(optimizer=torch.optim.SGD, optimizer_kwargs={"lr": 0.1}),
              loss=nn.MSELoss(), batch_size_per_device=2, gpu=False, configs=[B200Config(step_watchdog_s=30.0)])
    assert s.watchdog is not None and s.watchdog.beats == 0
    x, y = torch.randn(2, 4), torch.randn(2, 2)
    s.train_step(x, y)
    assert s.watchdog.beats >= 1
    s.watchdog.close()
    s2 = Stoke(model=nn.Linear(4, 2), verbose=False, optimizer=StokeOptimizer(optimizer=torch.optim.SGD, optimizer_kwargs={"lr": 0.1}),
               loss=nn.MSELoss(), batch_size_per_device=2, gpu=False)
    assert s2.watchdog is None


def test_swinir_grid_attention_and_physical_widths_match_the_roll_partition_blocks():
    """SwinIR's image-ordered blocks (window attention by addressing, projections on the whole grid) with activations
    zero-padded 60 -> 64 ("physical" widths) against the classic roll -> window_partition -> attention ->
    window_reverse -> roll blocks on the logical widths: same output, same gradients (fp64, CPU)."""
    import copy
    from pytorch_distributedtraining_b200 import models
    torch.manual_seed(0)
    m = models.SwinIR(upscale=2, in_chans=3, img_size=16, window_size=8, img_range=1., depths=[2, 2], embed_dim=60,
                      num_heads=[6, 6], mlp_ratio=2, upsampler="pixelshuffledirect", resi_connection="1conv",
                      drop_path_rate=0.0).double()
    for p in m.parameters():
        if p.dim() == 2 and p.shape[1] == 6:
            torch.nn.init.normal_(p, std=0.5)           # relative-position bias tables: make them matter
    legacy = copy.deepcopy(m)
    for mod in legacy.modules():
        if hasattr(mod, "legacy_block"):
            mod.legacy_block = True
    m.train(); legacy.train()
    m.phys_mode, legacy.phys_mode = True, False
    x = torch.rand(2, 3, 16, 24, dtype=torch.double)
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    y1, y2 = m(x1), legacy(x2)
    assert m._phys == 64 and legacy._phys is None
    torch.testing.assert_close(y1, y2, rtol=1e-10, atol=1e-12)
    y1.square().sum().backward(); y2.square().sum().backward()
    torch.testing.assert_close(x1.grad, x2.grad, rtol=1e-9, atol=1e-11)
    for (n, p), (_, q) in zip(m.named_parameters(), legacy.named_parameters()):
        torch.testing.assert_close(p.grad, q.grad, rtol=1e-8, atol=1e-10, msg=n)
    # the zero-padded parameter copies are cached per parameter version: an optimizer step must invalidate them
    for net in (m, legacy):
        torch.optim.SGD(net.parameters(), lr=1e-6).step()
        net.zero_grad()
    torch.testing.assert_close(m(x), legacy(x), rtol=1e-10, atol=1e-12)
    # ... and so must an update that bypasses the version counters (fused kernels, graph replays) once the epoch is bumped
    from pytorch_distributedtraining_b200.ops import fused
    with torch.no_grad():
        for net in (m, legacy):
            w = net.layers[0].residual_group.blocks[0].attn.qkv.weight
            w.data.view(-1)[:7] += 0.25          # .data: no version bump
    fused.bump_param_epoch()
    torch.testing.assert_close(m(x), legacy(x), rtol=1e-10, atol=1e-12)
    # switching the mode off again restores the logical widths everywhere
    m.phys_mode = False
    torch.testing.assert_close(m(x), legacy(x), rtol=1e-10, atol=1e-12)
    assert m.conv_first.phys_out is None and m.layers[0].residual_group.blocks[0].attn.qkv.phys_out is None


def test_padded_param_cache_semantics():
    """``ops.fused.padded_param``: one persistent zero-padded buffer per parameter, refreshed when the parameter's
    version or the global parameter epoch changed, sliced gradient, no cache for non-Parameter tensors."""
    from pytorch_distributedtraining_b200.ops import fused
    w = torch.nn.Parameter(torch.arange(6.0).view(2, 3))
    holder: dict = {}
    a = fused.padded_param(w, (4, 8), holder)
    assert a.shape == (4, 8) and torch.equal(a[:2, :3], w.detach()) and a[2:].abs().sum() == 0 and a[:, 3:].abs().sum() == 0
    buf = holder["buf"]
    b = fused.padded_param(w, (4, 8), holder)
    assert holder["buf"] is buf and b.data_ptr() == a.data_ptr()          # cache hit: same storage, nothing recomputed
    (a * torch.arange(32.0).view(4, 8)).sum().backward()
    torch.testing.assert_close(w.grad, torch.arange(32.0).view(4, 8)[:2, :3])
    with torch.no_grad():
        w.mul_(2.0)                                                       # version bump
    c = fused.padded_param(w, (4, 8), holder)
    assert torch.equal(c[:2, :3], w.detach())
    w.data[0, 0] = 100.0                                                  # no version bump: stale until the epoch moves
    assert fused.padded_param(w, (4, 8), holder)[0, 0] != 100.0
    fused.bump_param_epoch()
    assert fused.padded_param(w, (4, 8), holder)[0, 0] == 100.0
    assert fused.padded_param(w, (2, 3), holder) is w                     # nothing to pad
    t = w.detach() * 1.0                                                  # not a Parameter (e.g. an FSDP view): plain pad
    p = fused.padded_param(t, (4, 8), {})
    assert p.shape == (4, 8) and torch.equal(p[:2, :3], t)


def test_relative_position_bias_gather_gradient():
    """``_RelBias``: table[index] -> [heads, N, N]; its scatter-sum backward (index_add_ / sorted index_put_ in
    deterministic mode) against autograd's own gather backward."""
    from pytorch_distributedtraining_b200.models.swinir import _RelBias
    torch.manual_seed(0)
    table = torch.randn(225, 6, dtype=torch.double, requires_grad=True)
    index = torch.randint(0, 225, (64 * 64,))
    g = torch.randn(6, 64, 64, dtype=torch.double)
    ref = table[index].view(64, 64, -1).permute(2, 0, 1)
    (want,) = torch.autograd.grad(ref, table, g)
    for det in (False, True):
        torch.use_deterministic_algorithms(det)
        try:
            out = _RelBias.apply(table, index, 64)
            (got,) = torch.autograd.grad(out, table, g)
        finally:
            torch.use_deterministic_algorithms(False)
        assert torch.equal(out, ref)
        torch.testing.assert_close(got, want, rtol=1e-12, atol=1e-12)
