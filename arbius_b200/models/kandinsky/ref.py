"""Kandinsky 2.1 — eager PyTorch restatement of the networks behind the one model the reference actually enables
(reference: templates/kandinsky2.json, miner/src/config.json ``models.kandinsky2``; the container is the
``kasumi-1/Kandinsky-2`` fork at git aa5ee2f — not in tree, no weights offline: SURVEY §2.7 marks every dimension
"verify against real checkpoint key shapes"). Four networks:

  * ``TextEncoderXLMR``   XLM-R-Large style encoder (24 x 1024, post-LN, GELU) + mean pool + Linear -> 768
  * ``DiffusionPrior``    causal transformer (20 x 2048, 32 heads) that denoises the CLIP image embedding
  * ``KandinskyUNet``     guided-diffusion UNet: 384 x (1,2,3,4), 3 res blocks, scale-shift norm, resblock up/down,
                          attention at ds 2/4/8 with head dim 64 over [10 image-emb tokens + 77 text tokens + pixels],
                          in 4 / out 8 (learned sigma)
  * ``MoVQDecoder``       VQGAN decoder (128 x (1,2,2,4)) whose norms are modulated by the latent (SpatialNorm)

Parameter names follow the guided-diffusion layout the upstream package uses (``input_blocks.N.M``, ``middle_block``,
``output_blocks``, ``time_embed``, ``out``; ``in_layers``/``emb_layers``/``out_layers``/``skip_connection``; attention
``norm``/``qkv``/``encoder_kv``/``proj_out``).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..sd15.unet_ref import timestep_embedding


# --------------------------------------------------------------------------------------------- configs
@dataclass
class KUNetConfig:
    in_channels: int = 4
    out_channels: int = 8
    model_channels: int = 384
    channel_mult: Tuple[int, ...] = (1, 2, 3, 4)
    num_res_blocks: int = 3
    attention_ds: Tuple[int, ...] = (2, 4, 8)
    head_dim: int = 64
    text_dim: int = 1024          # XLM-R hidden size (token features)
    clip_dim: int = 768           # pooled text / image embedding size
    image_tokens: int = 10
    groups: int = 32

    @staticmethod
    def tiny() -> "KUNetConfig":
        return KUNetConfig(model_channels=64, channel_mult=(1, 2), num_res_blocks=1, attention_ds=(2,), head_dim=32,
                           text_dim=64, clip_dim=64, image_tokens=2, groups=8)


@dataclass
class PriorConfig:
    clip_dim: int = 768
    width: int = 2048
    layers: int = 20
    heads: int = 32
    text_len: int = 77
    timesteps: int = 1000

    @staticmethod
    def tiny() -> "PriorConfig":
        return PriorConfig(clip_dim=64, width=64, layers=2, heads=2)


@dataclass
class XlmrConfig:
    vocab_size: int = 250002
    hidden: int = 1024
    layers: int = 24
    heads: int = 16
    intermediate: int = 4096
    max_positions: int = 514
    out_dim: int = 768

    @staticmethod
    def tiny() -> "XlmrConfig":
        return XlmrConfig(vocab_size=1000, hidden=64, layers=2, heads=2, intermediate=128, max_positions=80, out_dim=64)


@dataclass
class MoVQConfig:
    z_channels: int = 4
    ch: int = 128
    ch_mult: Tuple[int, ...] = (1, 2, 2, 4)
    num_res_blocks: int = 2
    out_channels: int = 3
    groups: int = 32
    scaling: float = 1.0

    @staticmethod
    def tiny() -> "MoVQConfig":
        return MoVQConfig(ch=64, ch_mult=(1, 1), num_res_blocks=1, groups=8)


# --------------------------------------------------------------------------------------------- text encoder
class XlmrLayer(nn.Module):
    def __init__(self, c: XlmrConfig):
        super().__init__()
        self.heads = c.heads
        self.q, self.k, self.v = nn.Linear(c.hidden, c.hidden), nn.Linear(c.hidden, c.hidden), nn.Linear(c.hidden, c.hidden)
        self.o = nn.Linear(c.hidden, c.hidden)
        self.ln1 = nn.LayerNorm(c.hidden)
        self.fc1, self.fc2 = nn.Linear(c.hidden, c.intermediate), nn.Linear(c.intermediate, c.hidden)
        self.ln2 = nn.LayerNorm(c.hidden)

    def forward(self, x):
        b, l, d = x.shape
        sh = lambda t: t.view(b, l, self.heads, d // self.heads).transpose(1, 2)
        a = F.scaled_dot_product_attention(sh(self.q(x)), sh(self.k(x)), sh(self.v(x))).transpose(1, 2).reshape(b, l, d)
        x = self.ln1(x + self.o(a))                                     # post-LN (BERT/XLM-R)
        return self.ln2(x + self.fc2(F.gelu(self.fc1(x))))


class TextEncoderXLMR(nn.Module):
    def __init__(self, c: XlmrConfig = XlmrConfig()):
        super().__init__()
        self.cfg = c
        self.tok = nn.Embedding(c.vocab_size, c.hidden)
        self.pos = nn.Embedding(c.max_positions, c.hidden)
        self.ln_emb = nn.LayerNorm(c.hidden)
        self.layers = nn.ModuleList([XlmrLayer(c) for _ in range(c.layers)])
        self.proj = nn.Linear(c.hidden, c.out_dim)

    def forward(self, ids):
        """ids [B, L] -> (token features [B, L, hidden], pooled [B, out_dim])."""
        x = self.ln_emb(self.tok(ids) + self.pos(torch.arange(ids.shape[1], device=ids.device))[None])
        for layer in self.layers:
            x = layer(x)
        return x, self.proj(x.mean(dim=1))


# --------------------------------------------------------------------------------------------- prior
class PriorBlock(nn.Module):
    def __init__(self, c: PriorConfig):
        super().__init__()
        self.heads = c.heads
        self.ln_1 = nn.LayerNorm(c.width)
        self.c_qkv = nn.Linear(c.width, 3 * c.width)
        self.c_proj = nn.Linear(c.width, c.width)
        self.ln_2 = nn.LayerNorm(c.width)
        self.c_fc = nn.Linear(c.width, 4 * c.width)
        self.c_proj2 = nn.Linear(4 * c.width, c.width)

    def forward(self, x):
        b, l, d = x.shape
        q, k, v = self.c_qkv(self.ln_1(x)).chunk(3, dim=-1)
        sh = lambda t: t.view(b, l, self.heads, d // self.heads).transpose(1, 2)
        a = F.scaled_dot_product_attention(sh(q), sh(k), sh(v), is_causal=True).transpose(1, 2).reshape(b, l, d)
        x = x + self.c_proj(a)
        return x + self.c_proj2(F.gelu(self.c_fc(self.ln_2(x))))


class DiffusionPrior(nn.Module):
    """Sequence: [77 text tokens, pooled text emb, timestep emb, noisy image emb, learned query]; causal; the output
    at the query position is the predicted clean image embedding (x0 parameterisation)."""

    def __init__(self, c: PriorConfig = PriorConfig(), text_dim: int = 1024):
        super().__init__()
        self.cfg = c
        self.text_enc_proj = nn.Linear(text_dim, c.width)
        self.text_emb_proj = nn.Linear(c.clip_dim, c.width)
        self.time_embed = nn.Sequential(nn.Linear(c.width, c.width), nn.SiLU(), nn.Linear(c.width, c.width))
        self.x_proj = nn.Linear(c.clip_dim, c.width)
        self.query = nn.Parameter(torch.zeros(1, 1, c.width))
        self.pos = nn.Parameter(torch.zeros(1, c.text_len + 4, c.width))
        self.blocks = nn.ModuleList([PriorBlock(c) for _ in range(c.layers)])
        self.ln_final = nn.LayerNorm(c.width)
        self.out_proj = nn.Linear(c.width, c.clip_dim)

    def forward(self, x, t, text_tokens, text_emb):
        """x [B, clip_dim] noisy image emb, t [B], text_tokens [B, 77, text_dim], text_emb [B, clip_dim]."""
        b = x.shape[0]
        te = self.time_embed(timestep_embedding(t, self.cfg.width, False, 0.0).to(x.dtype))
        seq = torch.cat([self.text_enc_proj(text_tokens), self.text_emb_proj(text_emb)[:, None], te[:, None],
                         self.x_proj(x)[:, None], self.query.to(x.dtype).expand(b, -1, -1)], dim=1)
        h = seq + self.pos.to(x.dtype)
        for blk in self.blocks:
            h = blk(h)
        return self.out_proj(self.ln_final(h[:, -1]))


# --------------------------------------------------------------------------------------------- UNet
class KResBlock(nn.Module):
    def __init__(self, cin: int, cout: int, emb: int, groups: int, up: bool = False, down: bool = False):
        super().__init__()
        self.up, self.down = up, down
        self.in_layers = nn.Sequential(nn.GroupNorm(groups, cin), nn.SiLU(), nn.Conv2d(cin, cout, 3, padding=1))
        self.emb_layers = nn.Sequential(nn.SiLU(), nn.Linear(emb, 2 * cout))
        self.out_layers = nn.Sequential(nn.GroupNorm(groups, cout), nn.SiLU(), nn.Dropout(0.0), nn.Conv2d(cout, cout, 3, padding=1))
        self.skip_connection = nn.Conv2d(cin, cout, 1) if cin != cout else nn.Identity()

    def _resample(self, t):
        if self.up:
            return F.interpolate(t, scale_factor=2.0, mode="nearest")
        if self.down:
            return F.avg_pool2d(t, 2)
        return t

    def forward(self, x, emb):
        h = self.in_layers[1](self.in_layers[0](x))
        if self.up or self.down:
            h, x = self._resample(h), self._resample(x)
        h = self.in_layers[2](h)
        scale, shift = self.emb_layers(emb)[:, :, None, None].chunk(2, dim=1)
        h = self.out_layers[0](h) * (1 + scale) + shift
        h = self.out_layers[3](self.out_layers[1](h))
        return self.skip_connection(x) + h


class KAttentionBlock(nn.Module):
    def __init__(self, c: int, head_dim: int, enc_c: int, groups: int):
        super().__init__()
        self.heads = c // head_dim
        self.norm = nn.GroupNorm(groups, c)
        self.qkv = nn.Conv1d(c, 3 * c, 1)
        self.encoder_kv = nn.Conv1d(enc_c, 2 * c, 1)
        self.proj_out = nn.Conv1d(c, c, 1)

    def forward(self, x, enc):
        """x [B, C, H, W]; enc [B, enc_c, T] (image-embedding tokens then text tokens)."""
        b, c, hh, ww = x.shape
        xf = x.reshape(b, c, hh * ww)
        q, k, v = self.qkv(self.norm(xf)).chunk(3, dim=1)
        ek, ev = self.encoder_kv(enc).chunk(2, dim=1)
        k, v = torch.cat([ek, k], dim=-1), torch.cat([ev, v], dim=-1)
        sh = lambda t: t.reshape(b, self.heads, c // self.heads, -1).transpose(2, 3)
        a = F.scaled_dot_product_attention(sh(q), sh(k), sh(v)).transpose(2, 3).reshape(b, c, hh * ww)
        return (xf + self.proj_out(a)).reshape(b, c, hh, ww)


class KandinskyUNet(nn.Module):
    def __init__(self, c: KUNetConfig = KUNetConfig()):
        super().__init__()
        self.cfg = c
        mc, emb = c.model_channels, c.model_channels * 4
        self.time_embed = nn.Sequential(nn.Linear(mc, emb), nn.SiLU(), nn.Linear(emb, emb))
        self.ln_model_n = nn.LayerNorm(c.text_dim)
        self.to_model_dim_n = nn.Linear(c.text_dim, mc)
        self.proj_n = nn.Linear(c.clip_dim, emb)
        self.img_layer = nn.Linear(c.clip_dim, emb)
        self.clip_to_seq = nn.Linear(c.clip_dim, c.image_tokens * mc)
        g = c.groups
        self.input_blocks = nn.ModuleList([nn.ModuleList([nn.Conv2d(c.in_channels, mc, 3, padding=1)])])
        chans, ch, ds = [mc], mc, 1
        for level, mult in enumerate(c.channel_mult):
            for _ in range(c.num_res_blocks):
                layers: List[nn.Module] = [KResBlock(ch, mult * mc, emb, g)]
                ch = mult * mc
                if ds in c.attention_ds:
                    layers.append(KAttentionBlock(ch, c.head_dim, mc, g))
                self.input_blocks.append(nn.ModuleList(layers))
                chans.append(ch)
            if level != len(c.channel_mult) - 1:
                self.input_blocks.append(nn.ModuleList([KResBlock(ch, ch, emb, g, down=True)]))
                chans.append(ch)
                ds *= 2
        self.middle_block = nn.ModuleList([KResBlock(ch, ch, emb, g), KAttentionBlock(ch, c.head_dim, mc, g),
                                           KResBlock(ch, ch, emb, g)])
        self.output_blocks = nn.ModuleList()
        for level, mult in list(enumerate(c.channel_mult))[::-1]:
            for i in range(c.num_res_blocks + 1):
                ich = chans.pop()
                layers = [KResBlock(ch + ich, mult * mc, emb, g)]
                ch = mult * mc
                if ds in c.attention_ds:
                    layers.append(KAttentionBlock(ch, c.head_dim, mc, g))
                if level and i == c.num_res_blocks:
                    layers.append(KResBlock(ch, ch, emb, g, up=True))
                    ds //= 2
                self.output_blocks.append(nn.ModuleList(layers))
        self.out = nn.Sequential(nn.GroupNorm(g, ch), nn.SiLU(), nn.Conv2d(ch, c.out_channels, 3, padding=1))

    def conditioning(self, t, text_tokens, text_emb, image_emb):
        c = self.cfg
        emb = self.time_embed(timestep_embedding(t, c.model_channels, False, 0.0).to(text_emb.dtype))
        emb = emb + self.proj_n(text_emb) + self.img_layer(image_emb)
        seq = self.to_model_dim_n(self.ln_model_n(text_tokens))
        img = self.clip_to_seq(image_emb).reshape(t.shape[0], c.image_tokens, c.model_channels)
        return emb, torch.cat([img, seq], dim=1).transpose(1, 2)            # [B, mc, 10 + 77]

    def forward(self, x, t, text_tokens, text_emb, image_emb):
        emb, enc = self.conditioning(t, text_tokens, text_emb, image_emb)

        def run(layers, h):
            for m in layers:
                h = m(h, emb) if isinstance(m, KResBlock) else (m(h, enc) if isinstance(m, KAttentionBlock) else m(h))
            return h
        hs, h = [], x
        for blk in self.input_blocks:
            h = run(blk, h)
            hs.append(h)
        h = run(self.middle_block, h)
        for blk in self.output_blocks:
            h = run(blk, torch.cat([h, hs.pop()], dim=1))
        return self.out(h)


# --------------------------------------------------------------------------------------------- MoVQ decoder
class SpatialNorm(nn.Module):
    def __init__(self, c: int, zc: int, groups: int):
        super().__init__()
        self.norm_layer = nn.GroupNorm(groups, c, eps=1e-6)
        self.conv_y = nn.Conv2d(zc, c, 1)
        self.conv_b = nn.Conv2d(zc, c, 1)

    def forward(self, f, zq):
        zq = F.interpolate(zq, size=f.shape[-2:], mode="nearest")
        return self.norm_layer(f) * self.conv_y(zq) + self.conv_b(zq)


class MoVQResBlock(nn.Module):
    def __init__(self, cin: int, cout: int, zc: int, groups: int):
        super().__init__()
        self.norm1, self.conv1 = SpatialNorm(cin, zc, groups), nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2, self.conv2 = SpatialNorm(cout, zc, groups), nn.Conv2d(cout, cout, 3, padding=1)
        self.nin_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, zq):
        h = self.conv1(F.silu(self.norm1(x, zq)))
        h = self.conv2(F.silu(self.norm2(h, zq)))
        return (self.nin_shortcut(x) if self.nin_shortcut is not None else x) + h


class MoVQAttn(nn.Module):
    def __init__(self, c: int, zc: int, groups: int):
        super().__init__()
        self.norm = SpatialNorm(c, zc, groups)
        self.q, self.k, self.v, self.proj_out = (nn.Conv2d(c, c, 1) for _ in range(4))

    def forward(self, x, zq):
        b, c, h, w = x.shape
        n = self.norm(x, zq)
        sh = lambda t: t.reshape(b, 1, c, h * w).transpose(2, 3)
        a = F.scaled_dot_product_attention(sh(self.q(n)), sh(self.k(n)), sh(self.v(n))).transpose(2, 3).reshape(b, c, h, w)
        return x + self.proj_out(a)


class MoVQDecoder(nn.Module):
    def __init__(self, c: MoVQConfig = MoVQConfig()):
        super().__init__()
        self.cfg = c
        zc, g = c.z_channels, c.groups
        bi = c.ch * c.ch_mult[-1]
        self.post_quant_conv = nn.Conv2d(zc, zc, 1)
        self.conv_in = nn.Conv2d(zc, bi, 3, padding=1)
        self.mid_block_1, self.mid_attn_1, self.mid_block_2 = MoVQResBlock(bi, bi, zc, g), MoVQAttn(bi, zc, g), MoVQResBlock(bi, bi, zc, g)
        self.up = nn.ModuleList()
        for lvl in reversed(range(len(c.ch_mult))):
            bo = c.ch * c.ch_mult[lvl]
            blocks = nn.ModuleList()
            for _ in range(c.num_res_blocks + 1):
                blocks.append(MoVQResBlock(bi, bo, zc, g))
                bi = bo
            stage = nn.Module()
            stage.block = blocks
            stage.upsample = nn.Conv2d(bi, bi, 3, padding=1) if lvl != 0 else None
            self.up.append(stage)                                        # ordered from lowest resolution up
        self.norm_out = SpatialNorm(bi, zc, g)
        self.conv_out = nn.Conv2d(bi, c.out_channels, 3, padding=1)

    def forward(self, z):
        zq = self.post_quant_conv(z / self.cfg.scaling)
        h = self.conv_in(zq)
        h = self.mid_block_2(self.mid_attn_1(self.mid_block_1(h, zq), zq), zq)
        for stage in self.up:
            for blk in stage.block:
                h = blk(h, zq)
            if stage.upsample is not None:
                h = stage.upsample(F.interpolate(h, scale_factor=2.0, mode="nearest"))
        return self.conv_out(F.silu(self.norm_out(h, zq)))
