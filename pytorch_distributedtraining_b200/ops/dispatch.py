"""Which implementation serves a call: the sm_100a kernel or the stock torch op.

Policy: on a CUDA tensor the custom kernel is used whenever the (dtype, layout,
shape) is inside the kernel's envelope and the feature flag is on; the envelope is
stated here, in one place, so it can be read and tested.  Flags can be flipped with
``B200DP_DISABLE=conv,gemm,...`` for A/B measurements.
"""
from __future__ import annotations

import os
from typing import Any

import torch

_DISABLED = {s.strip() for s in os.environ.get("B200DP_DISABLE", "").split(",") if s.strip()}
_DETERMINISTIC = os.environ.get("B200DP_DETERMINISTIC", "0") not in ("", "0")
if _DETERMINISTIC:
    # run-to-run reproducible kernels only.  Three places accumulate with float atomics: the conv -> BatchNorm statistics
    # epilogue (BatchNorm goes back to its own two-stage, fixed-order reduction), the first-layer weight gradient (the layer
    # goes back to the library) and the bias gradient of window attention (see ``deterministic()`` in ops.attention)
    _DISABLED.add("convstats")
    _DISABLED.add("conv_first")
# measured on one GPU only so far (multi-rank NCCL runs without the peer engine keep the per-parameter clip loop unless
# B200DP_ENABLE=oss_fused_clip_multirank asks for the fused norm + one all-reduce)
if "oss_fused_clip_multirank" not in {s.strip() for s in os.environ.get("B200DP_ENABLE", "").split(",")}:
    _DISABLED.add("oss_fused_clip_multirank")
_HALF = (torch.float16, torch.bfloat16)


def deterministic() -> bool:
    """``B200DP_DETERMINISTIC=1`` or ``torch.use_deterministic_algorithms(True)``."""
    return _DETERMINISTIC or torch.are_deterministic_algorithms_enabled()


def enabled(feature: str) -> bool:
    return feature not in _DISABLED


def disable(feature: str) -> None:
    _DISABLED.add(feature)


def enable(feature: str) -> None:
    _DISABLED.discard(feature)


def _lib_ok() -> bool:
    from . import library_available
    return library_available()


def _autocast_dtype(x: torch.Tensor):
    if torch.is_autocast_enabled():
        return torch.get_autocast_dtype("cuda")
    return x.dtype


def bn_supported(mod: Any, x: torch.Tensor) -> bool:
    """NHWC (channels-last) fp16/bf16/fp32 activations, C % 8 == 0, affine, training or eval."""
    if not (enabled("bn") and _lib_ok()):
        return False
    if x.dim() != 4 or not mod.affine or not mod.track_running_stats:
        return False
    if x.dtype not in _HALF:
        return False
    if x.shape[1] % 8 != 0:
        return False
    return x.is_contiguous(memory_format=torch.channels_last)


def ln_supported(mod: Any, x: torch.Tensor) -> bool:
    if not (enabled("ln") and _lib_ok()):
        return False
    if len(mod.normalized_shape) != 1 or mod.weight is None or mod.bias is None:
        return False
    d = mod.normalized_shape[0]
    # multiples of 8 up to 8192; narrow rows (SwinIR-S: 60) on the 4-element-vector kernels up to 256
    return x.dtype in _HALF + (torch.float32,) and ((d % 8 == 0 and d <= 8192) or (d % 4 == 0 and d <= 256)) and x.is_contiguous()


def ln_phys_supported(mod: Any, x: torch.Tensor) -> bool:
    """LayerNorm over the first d elements of zero-padded rows (``ops.layers.LayerNorm.phys``)."""
    if not (enabled("ln") and _lib_ok()):
        return False
    if len(mod.normalized_shape) != 1 or mod.weight is None or mod.bias is None:
        return False
    d, p = mod.normalized_shape[0], x.shape[-1]
    return (x.dtype in _HALF + (torch.float32,) and d % 4 == 0 and p % 4 == 0 and d <= p <= 256 and x.is_contiguous())


def linear_phys_supported(w: torch.Tensor, x: torch.Tensor) -> bool:
    """A linear whose (already zero-padded) weight is inside the tcgen05 GEMM envelope."""
    if not (enabled("gemm") and _lib_ok()):
        return False
    from . import gemm
    if not gemm.available():
        return False
    return _autocast_dtype(x) in _HALF and w.shape[0] % 8 == 0 and w.shape[1] % 8 == 0 and x.shape[-1] == w.shape[1]


def conv_phys_supported(mod: Any, x: torch.Tensor, cin: int, cout: int) -> bool:
    """A convolution with zero-padded filters (``ops.layers.Conv2d.phys_in`` / ``phys_out``) inside the implicit-GEMM
    envelope: NHWC 16-bit compute, Cin % 64 == 0 (first layers: Cin <= 4), Cout % 8 == 0."""
    if not (enabled("conv") and _lib_ok()):
        return False
    from . import conv
    if not conv.available():
        return False
    dt = _autocast_dtype(x)
    if dt not in _HALF or mod.groups != 1 or mod.dilation != (1, 1) or mod.padding_mode != "zeros" or x.dim() != 4:
        return False
    if x.shape[1] != cin or cout % 8 != 0:
        return False
    sq = (mod.stride[0] == mod.stride[1] and mod.stride[0] in (1, 2) and mod.padding[0] == mod.padding[1]
          and isinstance(mod.padding[0], int) and x.shape[2] + 2 * mod.padding[0] >= mod.kernel_size[0])
    if cin <= 4:
        return (enabled("conv_first") and sq and mod.kernel_size[0] <= 7 and mod.kernel_size[1] <= 7 and cout <= 64)
    if mod.kernel_size == (1, 1) and mod.stride == (1, 1) and mod.padding == (0, 0):
        return cin % 8 == 0
    return sq and cin % 64 == 0 and x.is_contiguous(memory_format=torch.channels_last)


def linear_supported(mod: Any, x: torch.Tensor) -> bool:
    """tcgen05 GEMM envelope: fp16/bf16 compute (directly or under autocast), in/out features % 8 == 0
    (16-byte TMA row pitches); ragged M/N/K tile edges are handled in the kernel."""
    if not (enabled("gemm") and _lib_ok()):
        return False
    from . import gemm
    if not gemm.available():
        return False
    dt = _autocast_dtype(x)
    if dt not in _HALF:
        return False
    return mod.in_features % 8 == 0 and mod.out_features % 8 == 0


def linear_padded_supported(mod: Any, x: torch.Tensor) -> bool:
    """Widths that are not multiples of 8 (the reference's SwinIR-S: embed_dim 60, Stoke-DDP.py:206-208) still run on
    the tcgen05 GEMM: ``ops.layers.Linear`` zero-pads the activation's K and the weight's K / N up to the next
    multiple of 8 (16-byte TMA rows) and slices the result; 1-byte-per-element copies against keeping every linear of
    that model on the library."""
    if not (enabled("gemm") and enabled("pad") and _lib_ok()):
        return False
    from . import gemm
    if not gemm.available():
        return False
    return _autocast_dtype(x) in _HALF and mod.in_features >= 8 and mod.out_features >= 8


def conv_padded_supported(mod: Any, x: torch.Tensor) -> bool:
    """Convolutions whose channel counts miss the implicit-GEMM envelope by padding only (e.g. 60 -> 64): input
    channels are zero-padded to a multiple of 64, output channels to a multiple of 8 (64 when the input gradient is
    needed).  First-layer convolutions (Cin < 32) are excluded: padding 3 channels to 64 would move 21x the bytes."""
    if not (enabled("conv") and enabled("pad") and _lib_ok()):
        return False
    from . import conv
    if not conv.available():
        return False
    dt = _autocast_dtype(x)
    if dt not in _HALF or mod.groups != 1 or mod.dilation != (1, 1) or mod.padding_mode != "zeros":
        return False
    if x.dim() != 4 or mod.out_channels < 8:
        return False
    if mod.in_channels <= 4:
        # first layers whose Cout is not a multiple of 8 (SwinIR conv_first: 3 -> 60): pad the filters' Cout only
        return (enabled("conv_first") and mod.kernel_size[0] <= 7 and mod.kernel_size[1] <= 7 and mod.out_channels <= 64
                and mod.stride[0] == mod.stride[1] and mod.stride[0] in (1, 2)
                and mod.padding[0] == mod.padding[1] and isinstance(mod.padding[0], int))
    if mod.in_channels < 32:
        return False
    if mod.kernel_size == (1, 1) and mod.stride == (1, 1) and mod.padding == (0, 0):
        return False            # 1x1: handled as a (padded) linear
    return (mod.stride[0] == mod.stride[1] and mod.stride[0] in (1, 2) and mod.padding[0] == mod.padding[1]
            and isinstance(mod.padding[0], int) and x.shape[2] + 2 * mod.padding[0] >= mod.kernel_size[0])


def conv_supported(mod: Any, x: torch.Tensor) -> bool:
    """Implicit-GEMM envelope: NHWC fp16/bf16, groups == 1, dilation == 1, Cin % 64 == 0 (any multiple of 8 for 1x1 / stride 1), Cout % 8 == 0 — `ops.conv.shape_supported` is the check."""
    if not (enabled("conv") and _lib_ok()):
        return False
    from . import conv
    if not conv.available():
        return False
    dt = _autocast_dtype(x)
    if dt not in _HALF or mod.groups != 1 or mod.dilation != (1, 1) or mod.padding_mode != "zeros":
        return False
    return conv.shape_supported(mod, x)
