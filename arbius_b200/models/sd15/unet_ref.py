"""Stable Diffusion 1.5 UNet (Anything V3) — eager PyTorch restatement.

This is the numerical oracle (fp32 on CPU) and the timed "stock library kernels"
baseline (fp16 on GPU: cuDNN conv, cuBLAS, ATen GroupNorm, SDPA) that BASELINE.md §3
defines as the stand-in for the reference's Cog container
(reference: templates/anythingv3.json ``meta.docker``; the container code is not in tree,
SURVEY §2.7). Parameter names follow the diffusers ``UNet2DConditionModel`` checkpoint
layout so upstream safetensors load with ``load_state_dict`` unchanged.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class UNetConfig:
    in_channels: int = 4
    out_channels: int = 4
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    layers_per_block: int = 2
    attention_levels: Tuple[bool, ...] = (True, True, True, False)   # CrossAttnDown x3 + Down
    attention_head_dim: int = 8          # SD-1.x: this is the NUMBER of heads
    cross_attention_dim: int = 768
    norm_groups: int = 32
    norm_eps: float = 1e-5
    flip_sin_to_cos: bool = True
    freq_shift: float = 0.0

    @staticmethod
    def tiny() -> "UNetConfig":
        return UNetConfig(block_out_channels=(64, 128), layers_per_block=1, attention_levels=(True, False),
                          attention_head_dim=2, cross_attention_dim=64, norm_groups=8)


def timestep_embedding(t: torch.Tensor, dim: int, flip_sin_to_cos: bool = True, freq_shift: float = 0.0,
                       max_period: float = 10000.0) -> torch.Tensor:
    half = dim // 2
    exponent = -math.log(max_period) * torch.arange(half, dtype=torch.float32, device=t.device) / (half - freq_shift)
    emb = t.float()[:, None] * torch.exp(exponent)[None]
    emb = torch.cat([emb.sin(), emb.cos()], dim=-1)
    if flip_sin_to_cos:
        emb = torch.cat([emb[:, half:], emb[:, :half]], dim=-1)
    return emb


class TimestepEmbedding(nn.Module):
    def __init__(self, cin: int, dim: int):
        super().__init__()
        self.linear_1 = nn.Linear(cin, dim)
        self.linear_2 = nn.Linear(dim, dim)

    def forward(self, x):
        return self.linear_2(F.silu(self.linear_1(x)))


class ResnetBlock2D(nn.Module):
    def __init__(self, cin: int, cout: int, temb_dim: Optional[int], groups: int, eps: float):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_dim, cout) if temb_dim else None
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, temb=None):
        h = self.conv1(F.silu(self.norm1(x)))
        if self.time_emb_proj is not None and temb is not None:
            h = h + self.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return x + h


class Attention(nn.Module):
    def __init__(self, dim: int, heads: int, ctx_dim: Optional[int] = None, bias: bool = False):
        super().__init__()
        self.heads = heads
        ctx_dim = ctx_dim or dim
        self.to_q = nn.Linear(dim, dim, bias=bias)
        self.to_k = nn.Linear(ctx_dim, dim, bias=bias)
        self.to_v = nn.Linear(ctx_dim, dim, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(dim, dim)])

    def forward(self, x, ctx=None):
        ctx = x if ctx is None else ctx
        b, l, c = x.shape
        h = self.heads
        q = self.to_q(x).view(b, l, h, c // h).transpose(1, 2)
        k = self.to_k(ctx).view(b, ctx.shape[1], h, c // h).transpose(1, 2)
        v = self.to_v(ctx).view(b, ctx.shape[1], h, c // h).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v)
        return self.to_out[0](o.transpose(1, 2).reshape(b, l, c))


class GEGLU(nn.Module):
    def __init__(self, dim: int, inner: int):
        super().__init__()
        self.proj = nn.Linear(dim, inner * 2)

    def forward(self, x):
        h, gate = self.proj(x).chunk(2, dim=-1)
        return h * F.gelu(gate)


class FeedForward(nn.Module):
    def __init__(self, dim: int, mult: int = 4):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * mult), nn.Identity(), nn.Linear(dim * mult, dim)])

    def forward(self, x):
        return self.net[2](self.net[0](x))


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim: int, heads: int, ctx_dim: Optional[int]):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, heads, ctx_dim)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)

    def forward(self, x, ctx):
        x = x + self.attn1(self.norm1(x))
        x = x + self.attn2(self.norm2(x), ctx)
        return x + self.ff(self.norm3(x))


class Transformer2DModel(nn.Module):
    """SD-1.x variant: 1x1-conv proj_in / proj_out (use_linear_projection = False)."""

    def __init__(self, dim: int, heads: int, ctx_dim: int, groups: int):
        super().__init__()
        self.norm = nn.GroupNorm(groups, dim, eps=1e-6)
        self.proj_in = nn.Conv2d(dim, dim, 1)
        self.transformer_blocks = nn.ModuleList([BasicTransformerBlock(dim, heads, ctx_dim)])
        self.proj_out = nn.Conv2d(dim, dim, 1)

    def forward(self, x, ctx):
        b, c, h, w = x.shape
        y = self.proj_in(self.norm(x)).permute(0, 2, 3, 1).reshape(b, h * w, c)
        for blk in self.transformer_blocks:
            y = blk(y, ctx)
        y = y.reshape(b, h, w, c).permute(0, 3, 1, 2)
        return self.proj_out(y) + x


class Downsample2D(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, stride=2, padding=1)

    def forward(self, x):
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class DownBlock(nn.Module):
    def __init__(self, cfg: UNetConfig, cin: int, cout: int, temb: int, attn: bool, down: bool):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, temb, cfg.norm_groups, cfg.norm_eps)
                                      for i in range(cfg.layers_per_block)])
        self.attentions = nn.ModuleList([Transformer2DModel(cout, cfg.attention_head_dim, cfg.cross_attention_dim,
                                                            cfg.norm_groups)
                                         for _ in range(cfg.layers_per_block)]) if attn else None
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if down else None

    def forward(self, x, temb, ctx):
        outs = []
        for i, r in enumerate(self.resnets):
            x = r(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, ctx)
            outs.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
            outs.append(x)
        return x, outs


class UpBlock(nn.Module):
    def __init__(self, cfg: UNetConfig, cin: int, cout: int, prev: int, temb: int, attn: bool, up: bool):
        super().__init__()
        n = cfg.layers_per_block + 1
        res = []
        for i in range(n):
            skip = cin if i == n - 1 else cout
            inp = prev if i == 0 else cout
            res.append(ResnetBlock2D(inp + skip, cout, temb, cfg.norm_groups, cfg.norm_eps))
        self.resnets = nn.ModuleList(res)
        self.attentions = nn.ModuleList([Transformer2DModel(cout, cfg.attention_head_dim, cfg.cross_attention_dim,
                                                            cfg.norm_groups) for _ in range(n)]) if attn else None
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if up else None

    def forward(self, x, skips, temb, ctx):
        for i, r in enumerate(self.resnets):
            x = r(torch.cat([x, skips.pop()], dim=1), temb)
            if self.attentions is not None:
                x = self.attentions[i](x, ctx)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class MidBlock(nn.Module):
    def __init__(self, cfg: UNetConfig, c: int, temb: int):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, temb, cfg.norm_groups, cfg.norm_eps) for _ in range(2)])
        self.attentions = nn.ModuleList([Transformer2DModel(c, cfg.attention_head_dim, cfg.cross_attention_dim,
                                                            cfg.norm_groups)])

    def forward(self, x, temb, ctx):
        x = self.resnets[0](x, temb)
        x = self.attentions[0](x, ctx)
        return self.resnets[1](x, temb)


class UNet2DConditionModel(nn.Module):
    def __init__(self, cfg: UNetConfig = UNetConfig()):
        super().__init__()
        self.cfg = cfg
        ch = cfg.block_out_channels
        temb = ch[0] * 4
        self.conv_in = nn.Conv2d(cfg.in_channels, ch[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(ch[0], temb)
        downs, cin = [], ch[0]
        for i, cout in enumerate(ch):
            downs.append(DownBlock(cfg, cin, cout, temb, cfg.attention_levels[i], i < len(ch) - 1))
            cin = cout
        self.down_blocks = nn.ModuleList(downs)
        self.mid_block = MidBlock(cfg, ch[-1], temb)
        ups, rev = [], list(reversed(ch))
        rev_attn = list(reversed(cfg.attention_levels))
        prev = rev[0]
        for i, cout in enumerate(rev):
            cin_skip = rev[min(i + 1, len(ch) - 1)]
            ups.append(UpBlock(cfg, cin_skip, cout, prev, temb, rev_attn[i], i < len(ch) - 1))
            prev = cout
        self.up_blocks = nn.ModuleList(ups)
        self.conv_norm_out = nn.GroupNorm(cfg.norm_groups, ch[0], eps=cfg.norm_eps)
        self.conv_out = nn.Conv2d(ch[0], cfg.out_channels, 3, padding=1)

    def forward(self, x: torch.Tensor, t: torch.Tensor, ctx: torch.Tensor) -> torch.Tensor:
        """x [B,4,H,W], t [B] (or scalar), ctx [B,77,768] -> eps [B,4,H,W]."""
        if t.dim() == 0:
            t = t.expand(x.shape[0])
        temb = timestep_embedding(t, self.cfg.block_out_channels[0], self.cfg.flip_sin_to_cos, self.cfg.freq_shift)
        temb = self.time_embedding(temb.to(x.dtype))
        h = self.conv_in(x)
        skips = [h]
        for blk in self.down_blocks:
            h, outs = blk(h, temb, ctx)
            skips.extend(outs)
        h = self.mid_block(h, temb, ctx)
        for blk in self.up_blocks:
            h = blk(h, skips, temb, ctx)
        return self.conv_out(F.silu(self.conv_norm_out(h)))


def init_weights_(model: nn.Module, seed: int = 0) -> nn.Module:
    """Deterministic random init with unit-variance-preserving scales (no checkpoints offline)."""
    g = torch.Generator().manual_seed(seed)
    for name, p in sorted(model.named_parameters()):
        with torch.no_grad():
            if p.dim() >= 2:
                fan_in = p[0].numel()
                p.copy_(torch.randn(p.shape, generator=g) * (1.0 / math.sqrt(fan_in)))
            elif name.endswith("weight"):
                p.copy_(1.0 + 0.05 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
    return model
