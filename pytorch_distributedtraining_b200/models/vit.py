"""ViT-B/16 (and friends) for BASELINE.json config 4 (FSDP bf16).

Shapes and parameter count match torchvision's ``vit_b_16`` (86,567,656 parameters,
152 tensors).  Every transformer block is an FSDP wrap unit
(``fsdp_wrap_classes``): ≈7.1 M parameters → 14 MB in bf16, 1.8 MB per rank at 8 GPUs.

B200-first: linears go through ``ops.layers.Linear`` (tcgen05 GEMM with fused bias /
GELU epilogue on CUDA), LayerNorm through the fused kernel, attention through the
tcgen05 kernel pair of ``csrc/attn_sm100.cu`` (``ops.attention.attention``: q / k / v are read
from the packed projection, the backward writes the packed gradient).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import dispatch
from ..ops import gemm as _gemm
from ..ops.attention import attention, attention_reference, attention_supported
from ..ops.layers import LayerNorm, Linear

__all__ = ["VisionTransformer", "vit_b_16", "vit_s_16", "vit_tiny"]


class _PatchProj:
    """Shape carrier for ``dispatch.linear_supported``."""

    def __init__(self, in_features: int, out_features: int):
        self.in_features, self.out_features = in_features, out_features


class Attention(nn.Module):
    def __init__(self, dim: int, heads: int):
        super().__init__()
        self.heads = heads
        self.in_proj = Linear(dim, 3 * dim)
        self.out_proj = Linear(dim, dim)

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        qkv = self.in_proj(x)
        # tcgen05 attention over the packed projection (csrc/attn_sm100.cu); library kernel outside its envelope
        o = attention(qkv, self.heads) if attention_supported(qkv, self.heads) else attention_reference(qkv, self.heads)
        return self.out_proj(o, residual=residual)


class MLP(nn.Module):
    def __init__(self, dim: int, hidden: int):
        super().__init__()
        self.fc1 = Linear(dim, hidden, act="gelu")
        self.fc2 = Linear(hidden, dim)

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        if _gemm.mlp_supported(x, self.fc1.weight, self.fc2.weight):
            return _gemm.mlp(x, self.fc1.weight, self.fc1.bias, self.fc2.weight, self.fc2.bias, residual)
        return self.fc2(self.fc1(x), residual=residual)


class EncoderBlock(nn.Module):
    def __init__(self, dim: int, heads: int, mlp_dim: int):
        super().__init__()
        self.ln_1 = LayerNorm(dim, eps=1e-6)
        self.self_attention = Attention(dim, heads)
        self.ln_2 = LayerNorm(dim, eps=1e-6)
        self.mlp = MLP(dim, mlp_dim)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        # the residual stream rides through the kernels: LayerNorm hands its input back for the bypass
        # (forward_carry), the closing GEMM of each half adds it in its epilogue, and in the backward the
        # bypass gradient is folded into the LayerNorm backward — no stand-alone add kernels either way
        h, x = self.ln_1.forward_carry(x)
        x = self.self_attention(h, residual=x)
        h, x = self.ln_2.forward_carry(x)
        return self.mlp(h, residual=x)


class VisionTransformer(nn.Module):
    fsdp_wrap_classes = (EncoderBlock,)

    def __init__(self, image_size: int = 224, patch_size: int = 16, layers: int = 12, heads: int = 12,
                 hidden_dim: int = 768, mlp_dim: int = 3072, num_classes: int = 1000):
        super().__init__()
        self.patch_size, self.hidden_dim = patch_size, hidden_dim
        self.conv_proj = nn.Conv2d(3, hidden_dim, kernel_size=patch_size, stride=patch_size)
        seq = (image_size // patch_size) ** 2 + 1
        self.class_token = nn.Parameter(torch.zeros(1, 1, hidden_dim))
        self.pos_embedding = nn.Parameter(torch.empty(1, seq, hidden_dim).normal_(std=0.02))
        self.layers = nn.ModuleList([EncoderBlock(hidden_dim, heads, mlp_dim) for _ in range(layers)])
        self.ln = LayerNorm(hidden_dim, eps=1e-6)
        self.head = Linear(hidden_dim, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
        nn.init.zeros_(self.head.weight)

    def _embed_patches(self, x: torch.Tensor) -> torch.Tensor:
        """The stride-p p x p convolution of the patch embedding IS a matrix product over non-overlapping
        patches: one permute-copy of the image into ``[B*gh*gw, C*p*p]`` rows, then the tcgen05 GEMM with
        the convolution weight viewed as ``[D, C*p*p]`` (same parameter, same state-dict key)."""
        B, C, Hh, Ww = x.shape
        p = self.patch_size
        gh, gw = Hh // p, Ww // p
        w = self.conv_proj.weight
        probe = _PatchProj(w.shape[1] * p * p, w.shape[0])
        if not (x.is_cuda and dispatch.linear_supported(probe, x)):
            return self.conv_proj(x).flatten(2).transpose(1, 2)
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
        rows = x.to(dt).view(B, C, gh, p, gw, p).permute(0, 2, 4, 1, 3, 5).reshape(B * gh * gw, C * p * p)
        return _gemm.linear(rows, w.view(w.shape[0], -1), self.conv_proj.bias).view(B, gh * gw, w.shape[0])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B = x.shape[0]
        x = self._embed_patches(x)                                 # [B, S-1, D]
        x = torch.cat([self.class_token.expand(B, -1, -1).to(x.dtype), x], dim=1) + self.pos_embedding.to(x.dtype)
        for blk in self.layers:
            x = blk(x)
        # LayerNorm is per token: normalise only the class token that feeds the head
        return self.head(self.ln(x[:, 0]))


def vit_b_16(num_classes: int = 1000, image_size: int = 224) -> VisionTransformer:
    return VisionTransformer(image_size, 16, 12, 12, 768, 3072, num_classes)


def vit_s_16(num_classes: int = 1000, image_size: int = 224) -> VisionTransformer:
    return VisionTransformer(image_size, 16, 12, 6, 384, 1536, num_classes)


def vit_tiny(num_classes: int = 10, image_size: int = 32) -> VisionTransformer:
    """Test-sized model (CPU/gloo plumbing)."""
    return VisionTransformer(image_size, 8, 2, 2, 64, 128, num_classes)
