"""AutoencoderKL decoder (SD "better VAE": ch (128,256,512,512), 3 resnets per up block, single-head
512-ch mid attention, scale 0.18215) — eager oracle with diffusers parameter names
(``post_quant_conv``, ``decoder.*``). SURVEY §2.7 / §2.9 K14."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from .unet_ref import ResnetBlock2D


@dataclass
class VaeConfig:
    latent_channels: int = 4
    out_channels: int = 3
    block_out_channels: Tuple[int, ...] = (128, 256, 512, 512)
    layers_per_block: int = 2          # decoder uses layers_per_block + 1 resnets per up block
    norm_groups: int = 32
    scaling_factor: float = 0.18215

    @staticmethod
    def tiny() -> "VaeConfig":
        return VaeConfig(block_out_channels=(64, 64), layers_per_block=1, norm_groups=8)


class VaeAttention(nn.Module):
    def __init__(self, c: int, groups: int):
        super().__init__()
        self.group_norm = nn.GroupNorm(groups, c, eps=1e-6)
        self.to_q = nn.Linear(c, c)
        self.to_k = nn.Linear(c, c)
        self.to_v = nn.Linear(c, c)
        self.to_out = nn.ModuleList([nn.Linear(c, c)])

    def forward(self, x):
        b, c, h, w = x.shape
        y = self.group_norm(x).permute(0, 2, 3, 1).reshape(b, h * w, c)
        q, k, v = self.to_q(y)[:, None], self.to_k(y)[:, None], self.to_v(y)[:, None]
        o = F.scaled_dot_product_attention(q, k, v)[:, 0]
        o = self.to_out[0](o).reshape(b, h, w, c).permute(0, 3, 1, 2)
        return x + o


class VaeMidBlock(nn.Module):
    def __init__(self, c: int, groups: int):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, None, groups, 1e-6) for _ in range(2)])
        self.attentions = nn.ModuleList([VaeAttention(c, groups)])

    def forward(self, x):
        return self.resnets[1](self.attentions[0](self.resnets[0](x)))


class VaeUpsample(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class VaeUpBlock(nn.Module):
    def __init__(self, cin: int, cout: int, n: int, groups: int, up: bool):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, None, groups, 1e-6) for i in range(n)])
        self.upsamplers = nn.ModuleList([VaeUpsample(cout)]) if up else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        return self.upsamplers[0](x) if self.upsamplers is not None else x


class Decoder(nn.Module):
    def __init__(self, cfg: VaeConfig):
        super().__init__()
        ch = cfg.block_out_channels
        self.conv_in = nn.Conv2d(cfg.latent_channels, ch[-1], 3, padding=1)
        self.mid_block = VaeMidBlock(ch[-1], cfg.norm_groups)
        rev = list(reversed(ch))
        ups, prev = [], rev[0]
        for i, c in enumerate(rev):
            ups.append(VaeUpBlock(prev, c, cfg.layers_per_block + 1, cfg.norm_groups, i < len(ch) - 1))
            prev = c
        self.up_blocks = nn.ModuleList(ups)
        self.conv_norm_out = nn.GroupNorm(cfg.norm_groups, ch[0], eps=1e-6)
        self.conv_out = nn.Conv2d(ch[0], cfg.out_channels, 3, padding=1)

    def forward(self, z):
        h = self.mid_block(self.conv_in(z))
        for blk in self.up_blocks:
            h = blk(h)
        return self.conv_out(F.silu(self.conv_norm_out(h)))


class AutoencoderKLDecoder(nn.Module):
    """Decoder half only (the miner never encodes images for txt2img)."""

    def __init__(self, cfg: VaeConfig = VaeConfig()):
        super().__init__()
        self.cfg = cfg
        self.post_quant_conv = nn.Conv2d(cfg.latent_channels, cfg.latent_channels, 1)
        self.decoder = Decoder(cfg)

    def forward(self, latents: torch.Tensor) -> torch.Tensor:
        """latents [B,4,h,w] (scaled) -> image [B,3,8h,8w] in [-1,1]."""
        return self.decoder(self.post_quant_conv(latents / self.cfg.scaling_factor))
