"""UNet3DConditionModel (ModelScope text-to-video family: ZeroScope v2 XL, DAMO ms-1.7b) — eager oracle.

Layout and parameter names follow the diffusers ``UNet3DConditionModel`` checkpoints the reference's video
templates point at (reference: templates/zeroscopev2xl.json, templates/damo.json ``meta.docker``/``git``; the
container code itself is not in tree — SURVEY §2.7): every spatial stage is ResnetBlock2D -> TemporalConvLayer
-> Transformer2D (cross-attn on text) -> TransformerTemporalModel (two self-attentions over the frame axis),
plus ``transformer_in`` after conv_in. Spatial layers see frames as batch; temporal layers see pixels as batch.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..sd15.unet_ref import (Attention, Downsample2D, FeedForward, ResnetBlock2D, TimestepEmbedding, Upsample2D,
                             timestep_embedding)


@dataclass
class UNet3DConfig:
    in_channels: int = 4
    out_channels: int = 4
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    layers_per_block: int = 2
    attention_levels: Tuple[bool, ...] = (True, True, True, False)
    attention_head_dim: int = 64           # here it really is the head dim: heads = C / 64
    cross_attention_dim: int = 1024
    norm_groups: int = 32
    norm_eps: float = 1e-5

    @staticmethod
    def tiny() -> "UNet3DConfig":
        return UNet3DConfig(block_out_channels=(64, 128), layers_per_block=1, attention_levels=(True, False),
                            attention_head_dim=32, cross_attention_dim=64, norm_groups=8)


class TemporalConvLayer(nn.Module):
    """4 x [GroupNorm, SiLU, (Dropout), Conv3d (3,1,1)] with identity skip; conv4 starts at zero upstream."""

    def __init__(self, c: int, groups: int):
        super().__init__()
        def blk(first):
            layers = [nn.GroupNorm(groups, c), nn.SiLU()]
            if not first:
                layers.append(nn.Dropout(0.0))
            layers.append(nn.Conv3d(c, c, (3, 1, 1), padding=(1, 0, 0)))
            return nn.Sequential(*layers)
        self.conv1, self.conv2, self.conv3, self.conv4 = blk(True), blk(False), blk(False), blk(False)

    def forward(self, x, num_frames: int):
        bf, c, h, w = x.shape
        y = x.reshape(bf // num_frames, num_frames, c, h, w).permute(0, 2, 1, 3, 4)
        idt = y
        y = self.conv4(self.conv3(self.conv2(self.conv1(y))))
        y = idt + y
        return y.permute(0, 2, 1, 3, 4).reshape(bf, c, h, w)


class TransformerBlock(nn.Module):
    """BasicTransformerBlock; ``double_self`` makes attn2 a second self-attention (temporal transformer)."""

    def __init__(self, dim: int, heads: int, ctx_dim: Optional[int], double_self: bool = False):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, heads, None if double_self else ctx_dim)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)
        self.double_self = double_self

    def forward(self, x, ctx=None):
        x = x + self.attn1(self.norm1(x))
        x = x + self.attn2(self.norm2(x), None if self.double_self else ctx)
        return x + self.ff(self.norm3(x))


class Transformer2DLinear(nn.Module):
    """Spatial transformer with linear projections (use_linear_projection=True in the video UNets)."""

    def __init__(self, dim: int, heads: int, ctx_dim: int, groups: int):
        super().__init__()
        self.norm = nn.GroupNorm(groups, dim, eps=1e-6)
        self.proj_in = nn.Linear(dim, dim)
        self.transformer_blocks = nn.ModuleList([TransformerBlock(dim, heads, ctx_dim)])
        self.proj_out = nn.Linear(dim, dim)

    def forward(self, x, ctx):
        b, c, h, w = x.shape
        y = self.norm(x).permute(0, 2, 3, 1).reshape(b, h * w, c)
        y = self.proj_in(y)
        for blk in self.transformer_blocks:
            y = blk(y, ctx)
        y = self.proj_out(y).reshape(b, h, w, c).permute(0, 3, 1, 2)
        return y + x


class TransformerTemporalModel(nn.Module):
    def __init__(self, dim: int, heads: int, groups: int, head_dim: Optional[int] = None):
        super().__init__()
        inner = heads * head_dim if head_dim else dim        # transformer_in: 8 heads x 64 = 512 on 320 channels
        self.norm = nn.GroupNorm(groups, dim, eps=1e-6)
        self.proj_in = nn.Linear(dim, inner)
        self.transformer_blocks = nn.ModuleList([TransformerBlock(inner, heads, None, double_self=True)])
        self.proj_out = nn.Linear(inner, dim)

    def forward(self, x, num_frames: int):
        bf, c, h, w = x.shape
        b = bf // num_frames
        res = x
        y = x.reshape(b, num_frames, c, h, w).permute(0, 2, 1, 3, 4)          # [B, C, F, H, W]
        y = self.norm(y)
        y = y.permute(0, 3, 4, 2, 1).reshape(b * h * w, num_frames, c)        # pixels as batch
        y = self.proj_in(y)
        for blk in self.transformer_blocks:
            y = blk(y)
        y = self.proj_out(y)
        y = y.reshape(b, h, w, num_frames, c).permute(0, 3, 4, 1, 2).reshape(bf, c, h, w)
        return y + res


class DownBlock3D(nn.Module):
    def __init__(self, cfg, cin, cout, temb, attn, down):
        super().__init__()
        n, g = cfg.layers_per_block, cfg.norm_groups
        heads = cout // cfg.attention_head_dim
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, temb, g, cfg.norm_eps) for i in range(n)])
        self.temp_convs = nn.ModuleList([TemporalConvLayer(cout, g) for _ in range(n)])
        self.attentions = nn.ModuleList([Transformer2DLinear(cout, heads, cfg.cross_attention_dim, g) for _ in range(n)]) if attn else None
        self.temp_attentions = nn.ModuleList([TransformerTemporalModel(cout, heads, g) for _ in range(n)]) if attn else None
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if down else None

    def forward(self, x, temb, ctx, nf):
        outs = []
        for i, r in enumerate(self.resnets):
            x = self.temp_convs[i](r(x, temb), nf)
            if self.attentions is not None:
                x = self.temp_attentions[i](self.attentions[i](x, ctx), nf)
            outs.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
            outs.append(x)
        return x, outs


class UpBlock3D(nn.Module):
    def __init__(self, cfg, cin, cout, prev, temb, attn, up):
        super().__init__()
        n, g = cfg.layers_per_block + 1, cfg.norm_groups
        heads = cout // cfg.attention_head_dim
        res = []
        for i in range(n):
            skip = cin if i == n - 1 else cout
            inp = prev if i == 0 else cout
            res.append(ResnetBlock2D(inp + skip, cout, temb, g, cfg.norm_eps))
        self.resnets = nn.ModuleList(res)
        self.temp_convs = nn.ModuleList([TemporalConvLayer(cout, g) for _ in range(n)])
        self.attentions = nn.ModuleList([Transformer2DLinear(cout, heads, cfg.cross_attention_dim, g) for _ in range(n)]) if attn else None
        self.temp_attentions = nn.ModuleList([TransformerTemporalModel(cout, heads, g) for _ in range(n)]) if attn else None
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if up else None

    def forward(self, x, skips, temb, ctx, nf):
        for i, r in enumerate(self.resnets):
            x = self.temp_convs[i](r(torch.cat([x, skips.pop()], dim=1), temb), nf)
            if self.attentions is not None:
                x = self.temp_attentions[i](self.attentions[i](x, ctx), nf)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class MidBlock3D(nn.Module):
    def __init__(self, cfg, c, temb):
        super().__init__()
        g, heads = cfg.norm_groups, c // cfg.attention_head_dim
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, temb, g, cfg.norm_eps) for _ in range(2)])
        self.temp_convs = nn.ModuleList([TemporalConvLayer(c, g) for _ in range(2)])
        self.attentions = nn.ModuleList([Transformer2DLinear(c, heads, cfg.cross_attention_dim, g)])
        self.temp_attentions = nn.ModuleList([TransformerTemporalModel(c, heads, g)])

    def forward(self, x, temb, ctx, nf):
        x = self.temp_convs[0](self.resnets[0](x, temb), nf)
        x = self.temp_attentions[0](self.attentions[0](x, ctx), nf)
        return self.temp_convs[1](self.resnets[1](x, temb), nf)


class UNet3DConditionModel(nn.Module):
    def __init__(self, cfg: UNet3DConfig = UNet3DConfig()):
        super().__init__()
        self.cfg = cfg
        ch = cfg.block_out_channels
        temb = ch[0] * 4
        self.conv_in = nn.Conv2d(cfg.in_channels, ch[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(ch[0], temb)
        self.transformer_in = TransformerTemporalModel(ch[0], 8, cfg.norm_groups, cfg.attention_head_dim)
        downs, cin = [], ch[0]
        for i, cout in enumerate(ch):
            downs.append(DownBlock3D(cfg, cin, cout, temb, cfg.attention_levels[i], i < len(ch) - 1))
            cin = cout
        self.down_blocks = nn.ModuleList(downs)
        self.mid_block = MidBlock3D(cfg, ch[-1], temb)
        ups, rev, rev_attn, prev = [], list(reversed(ch)), list(reversed(cfg.attention_levels)), ch[-1]
        for i, cout in enumerate(rev):
            cin_skip = rev[min(i + 1, len(ch) - 1)]
            ups.append(UpBlock3D(cfg, cin_skip, cout, prev, temb, rev_attn[i], i < len(ch) - 1))
            prev = cout
        self.up_blocks = nn.ModuleList(ups)
        self.conv_norm_out = nn.GroupNorm(cfg.norm_groups, ch[0], eps=cfg.norm_eps)
        self.conv_out = nn.Conv2d(ch[0], cfg.out_channels, 3, padding=1)

    def forward(self, x: torch.Tensor, t: torch.Tensor, ctx: torch.Tensor) -> torch.Tensor:
        """x [B, 4, F, H, W], t [B], ctx [B, 77, D] -> eps [B, 4, F, H, W]."""
        b, c, nf, h, w = x.shape
        if t.dim() == 0:
            t = t.expand(b)
        temb = self.time_embedding(timestep_embedding(t, self.cfg.block_out_channels[0], True, 0.0).to(x.dtype))
        temb = temb.repeat_interleave(nf, dim=0)
        ctx_f = ctx.repeat_interleave(nf, dim=0)
        y = x.permute(0, 2, 1, 3, 4).reshape(b * nf, c, h, w)
        y = self.conv_in(y)
        y = self.transformer_in(y, nf)
        skips = [y]
        for blk in self.down_blocks:
            y, outs = blk(y, temb, ctx_f, nf)
            skips.extend(outs)
        y = self.mid_block(y, temb, ctx_f, nf)
        for blk in self.up_blocks:
            y = blk(y, skips, temb, ctx_f, nf)
        y = self.conv_out(F.silu(self.conv_norm_out(y)))
        return y.reshape(b, nf, -1, h, w).permute(0, 2, 1, 3, 4)
