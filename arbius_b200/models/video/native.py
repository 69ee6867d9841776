"""UNet3D (ZeroScope v2 XL / DAMO) on the sm_100a kernels.

Activations are channels-last video tensors [B, F, H, W, C]. Because the token order is (b, f, y, x):
  * spatial layers (resnets, spatial transformers) see the same memory as [B*F, H, W, C] images;
  * temporal convs see it as the image [B, F, H*W, C] with a 3x1 tap pattern (implicit GEMM, no permute);
  * temporal attention tiles are assembled by 5-D TMA boxes (frames x pixels) straight from that layout.
So the frame<->pixel "transposes" of the eager model never materialise. Weights come from the same state dict
as the eager oracle (diffusers UNet3DConditionModel key layout) and are repacked once.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ... import ops
from ...ops.temporal import temporal_attn
from ..sd15.native import _Resnet, _get
from .unet3d_ref import UNet3DConfig

SD = Dict[str, torch.Tensor]


class _TemporalConv:
    def __init__(self, sd: SD, p: str, dtype, device, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.groups = groups
        self.stages = []
        for name, ci in (("conv1", 2), ("conv2", 3), ("conv3", 3), ("conv4", 3)):
            self.stages.append((g(f"{name}.0.weight"), g(f"{name}.0.bias"),
                                ops.pack_temporal_conv_weight(g(f"{name}.{ci}.weight")), g(f"{name}.{ci}.bias")))

    shard = None                       # a parallel.frame_shard.FrameShard when the frames are spread over GPUs

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        """x [B, F, H, W, C]"""
        b, f, h, w, c = x.shape
        y = x.view(b, f, h * w, c)
        idt = y
        if self.shard is not None:
            for i, (nw, nb, cw, cb) in enumerate(self.stages):
                pad = self.shard.groupnorm(y, nw, nb, self.groups, 1e-5, "silu", padded_parity=i & 1)
                self.shard.exchange_halos(pad, i & 1)
                y = ops.conv_temporal(pad, cw, cb, residual=idt if i == 3 else None, halo=True,
                                      frame_ranks=self.shard.world)
            return y.view(b, f, h, w, c)
        for i, (nw, nb, cw, cb) in enumerate(self.stages):
            y = ops.groupnorm(y, nw, nb, self.groups, 1e-5, "silu")           # stats over (F, H, W) per sample
            y = ops.conv_temporal(y, cw, cb, residual=idt if i == 3 else None)
        return y.view(b, f, h, w, c)


import threading

TemporalShardInfo = threading.local()      # .ranks: how many ranks the clip's frames are spread over in THIS thread's forward


class _Block:
    """BasicTransformerBlock weights (attn1 self, attn2 cross or second self, GEGLU feed-forward)."""

    def __init__(self, sd: SD, t: str, dtype, device, double_self: bool):
        g = lambda k: _get(sd, t + k, dtype, device)
        self.double_self = double_self
        self.ln1 = (g("norm1.weight"), g("norm1.bias"))
        self.ln2 = (g("norm2.weight"), g("norm2.bias"))
        self.ln3 = (g("norm3.weight"), g("norm3.bias"))
        self.wqkv1 = torch.cat([g("attn1.to_q.weight"), g("attn1.to_k.weight"), g("attn1.to_v.weight")], 0).contiguous()
        self.wo1, self.bo1 = g("attn1.to_out.0.weight"), g("attn1.to_out.0.bias")
        if double_self:
            self.wqkv2 = torch.cat([g("attn2.to_q.weight"), g("attn2.to_k.weight"), g("attn2.to_v.weight")], 0).contiguous()
        else:
            self.wq2 = g("attn2.to_q.weight")
            self.wkv2 = torch.cat([g("attn2.to_k.weight"), g("attn2.to_v.weight")], 0).contiguous()
        self.wo2, self.bo2 = g("attn2.to_out.0.weight"), g("attn2.to_out.0.bias")
        self.wff1, self.bff1 = ops.pack_geglu_weight(g("ff.net.0.proj.weight"), g("ff.net.0.proj.bias"))
        self.wff2, self.bff2 = g("ff.net.2.weight"), g("ff.net.2.bias")

    def feed_forward(self, t):
        n = ops.layernorm(t, *self.ln3)
        f = ops.gemm(n, self.wff1, self.bff1, geglu=True)
        return ops.gemm(f, self.wff2, self.bff2, residual=t)


class _SpatialTransformer:
    def __init__(self, sd: SD, p: str, dtype, device, heads: int, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.heads, self.groups = heads, groups
        self.nw, self.nb = g("norm.weight"), g("norm.bias")
        self.piw, self.pib = g("proj_in.weight"), g("proj_in.bias")
        self.pow, self.pob = g("proj_out.weight"), g("proj_out.bias")
        self.c = self.piw.shape[0]
        self.blk = _Block(sd, p + "transformer_blocks.0.", dtype, device, double_self=False)

    def context_kv(self, ctx):
        return ops.gemm(ctx, self.blk.wkv2)

    def __call__(self, x: torch.Tensor, kv: torch.Tensor) -> torch.Tensor:
        """x [B, F, H, W, C]; kv [B, 77, 2C] (shared by all frames of a sample)."""
        b, f, hh, ww, c = x.shape
        l, h, d = hh * ww, self.heads, c // self.heads
        B = self.blk
        xf = x.view(b * f, l, c)
        t = ops.groupnorm(xf, self.nw, self.nb, self.groups, 1e-6, None)        # per frame
        t = ops.gemm(t.view(b * f * l, c), self.piw, self.pib)
        n = ops.layernorm(t, *B.ln1)
        qkv = ops.gemm(n, B.wqkv1).view(b * f, l, 3 * c)
        q, k, v = (qkv[..., i * c:(i + 1) * c].view(b * f, l, h, d) for i in range(3))
        a = ops.flash_attn(q, k, v)
        t = ops.gemm(a.view(b * f * l, c), B.wo1, B.bo1, residual=t)
        n = ops.layernorm(t, *B.ln2)
        q = ops.gemm(n, B.wq2).view(b, f * l, h, d)                              # frames of a sample share its text
        lk = kv.shape[1]
        # a frame-sharded rank holds f/P of the sample's query rows: the kernel choice must be the whole clip's
        ranks = getattr(TemporalShardInfo, "ranks", 1)
        a = ops.flash_attn(q, kv[..., :c].view(b, lk, h, d), kv[..., c:].view(b, lk, h, d),
                           long_queries=ranks > 1 and f * l * ranks > 128)
        t = ops.gemm(a.view(b * f * l, c), B.wo2, B.bo2, residual=t)
        t = B.feed_forward(t)
        return ops.gemm(t, self.pow, self.pob, residual=x.view(b * f * l, c)).view(b, f, hh, ww, c)


class _TemporalTransformer:
    def __init__(self, sd: SD, p: str, dtype, device, heads: int, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.heads, self.groups = heads, groups
        self.nw, self.nb = g("norm.weight"), g("norm.bias")
        self.piw, self.pib = g("proj_in.weight"), g("proj_in.bias")
        self.pow, self.pob = g("proj_out.weight"), g("proj_out.bias")
        self.inner = self.piw.shape[0]
        self.blk = _Block(sd, p + "transformer_blocks.0.", dtype, device, double_self=True)

    shard = None                       # see _TemporalConv.shard

    def _self_attn(self, t, ln, wqkv, wo, bo, b, f, hw):
        c = self.inner
        n = ops.layernorm(t, *ln)
        if self.shard is not None:
            ops.gemm(n, wqkv, out=self.shard.qkv_buffer(b, hw, 3 * c))
            a = self.shard.temporal_attention(b, hw, self.heads, c)
            return ops.gemm(a, wo, bo, residual=t)
        qkv = ops.gemm(n, wqkv).view(b, f, hw, 3 * c)
        q, k, v = (qkv[..., i * c:(i + 1) * c] for i in range(3))
        a = temporal_attn(q, k, v, self.heads)
        return ops.gemm(a.view(b * f * hw, c), wo, bo, residual=t)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        b, f, hh, ww, c = x.shape
        hw = hh * ww
        B = self.blk
        if self.shard is not None:
            t = self.shard.groupnorm(x.view(b, f * hw, c), self.nw, self.nb, self.groups, 1e-6, None)
        else:
            t = ops.groupnorm(x.view(b, f * hw, c), self.nw, self.nb, self.groups, 1e-6, None)     # over (F, H, W)
        t = ops.gemm(t.view(b * f * hw, c), self.piw, self.pib)
        t = self._self_attn(t, B.ln1, B.wqkv1, B.wo1, B.bo1, b, f, hw)
        t = self._self_attn(t, B.ln2, B.wqkv2, B.wo2, B.bo2, b, f, hw)
        t = B.feed_forward(t)
        return ops.gemm(t, self.pow, self.pob, residual=x.view(b * f * hw, c)).view(b, f, hh, ww, c)


class NativeUNet3D:
    def __init__(self, sd: SD, cfg: UNet3DConfig = UNet3DConfig(), dtype=torch.float16, device="cuda"):
        self.cfg, self.dtype = cfg, dtype
        g = lambda k: _get(sd, k, dtype, device)
        ch, gr, hd = cfg.block_out_channels, cfg.norm_groups, cfg.attention_head_dim
        self.conv_in_w = ops.DirectConvWeight(g("conv_in.weight").permute(0, 2, 3, 1))
        self.conv_in_b = g("conv_in.bias")
        self.te = (g("time_embedding.linear_1.weight"), g("time_embedding.linear_1.bias"),
                   g("time_embedding.linear_2.weight"), g("time_embedding.linear_2.bias"))
        self.transformer_in = _TemporalTransformer(sd, "transformer_in.", dtype, device, 8, gr)
        mk_r = lambda p: _Resnet(sd, p, dtype, device, cfg.norm_eps, gr)
        mk_c = lambda p: _TemporalConv(sd, p, dtype, device, gr)
        mk_s = lambda p, c: _SpatialTransformer(sd, p, dtype, device, c // hd, gr)
        mk_t = lambda p, c: _TemporalTransformer(sd, p, dtype, device, c // hd, gr)

        def stage(prefix, n, c, attn):
            return {"res": [mk_r(f"{prefix}resnets.{j}.") for j in range(n)],
                    "tc": [mk_c(f"{prefix}temp_convs.{j}.") for j in range(n)],
                    "att": [mk_s(f"{prefix}attentions.{j}.", c) for j in range(n)] if attn else None,
                    "tatt": [mk_t(f"{prefix}temp_attentions.{j}.", c) for j in range(n)] if attn else None}
        self.down = []
        for i, c in enumerate(ch):
            blk = stage(f"down_blocks.{i}.", cfg.layers_per_block, c, cfg.attention_levels[i])
            blk["down"] = None
            if i < len(ch) - 1:
                p = f"down_blocks.{i}.downsamplers.0.conv."
                blk["down"] = (ops.pack_conv_weight(g(p + "weight")), g(p + "bias"))
            self.down.append(blk)
        c = ch[-1]
        self.mid = {"res": [mk_r("mid_block.resnets.0."), mk_r("mid_block.resnets.1.")],
                    "tc": [mk_c("mid_block.temp_convs.0."), mk_c("mid_block.temp_convs.1.")],
                    "att": mk_s("mid_block.attentions.0.", c), "tatt": mk_t("mid_block.temp_attentions.0.", c)}
        self.up = []
        rev, rev_attn = list(reversed(ch)), list(reversed(cfg.attention_levels))
        for i, c in enumerate(rev):
            blk = stage(f"up_blocks.{i}.", cfg.layers_per_block + 1, c, rev_attn[i])
            blk["up"] = None
            if i < len(ch) - 1:
                p = f"up_blocks.{i}.upsamplers.0.conv."
                blk["up"] = (ops.pack_conv_weight(g(p + "weight")), g(p + "bias"))
            self.up.append(blk)
        self.no_w, self.no_b = g("conv_norm_out.weight"), g("conv_norm_out.bias")
        self.conv_out_w = ops.DirectConvWeight(g("conv_out.weight").permute(0, 2, 3, 1))
        self.conv_out_b = g("conv_out.bias")
        res_all: List[_Resnet] = []
        for blk in self.down:
            res_all += blk["res"]
        res_all += self.mid["res"]
        for blk in self.up:
            res_all += blk["res"]
        off = 0
        for r in res_all:
            r.temb_off = off
            off += r.cout
        self.temb_w = torch.cat([r.tw for r in res_all], 0).contiguous()
        self.temb_b = torch.cat([r.tb for r in res_all], 0).contiguous()
        for r in res_all:
            r.tw = r.tb = None
        self.spatial: List[_SpatialTransformer] = []
        for blk in self.down:
            self.spatial += blk["att"] or []
        self.spatial.append(self.mid["att"])
        for blk in self.up:
            self.spatial += blk["att"] or []

    def precompute_context(self, ctx: torch.Tensor) -> List[torch.Tensor]:
        return [t.context_kv(ctx) for t in self.spatial]

    def set_frame_shard(self, shard) -> None:
        """Run the temporal layers across the ranks of ``shard`` (parallel/frame_shard.py); None: single GPU."""
        self.frame_ranks = shard.world if shard is not None else 1
        mods = [self.transformer_in, self.mid["tatt"]] + self.mid["tc"]
        for blk in self.down + self.up:
            mods += blk["tc"] + (blk["tatt"] or [])
        for m in mods:
            m.shard = shard

    def time_projections(self, t: torch.Tensor, frames: int) -> torch.Tensor:
        """t [B] -> [B*F, sum(Cout)]: one row per (sample, frame) image, as the spatial resnets index it."""
        e = ops.timestep_embedding(t, self.cfg.block_out_channels[0], self.dtype, True, 0.0)
        e = ops.gemm(e, self.te[0], self.te[1], act="silu", splitk=1)
        e = ops.gemm(e, self.te[2], self.te[3], act="silu", splitk=1)
        proj = ops.gemm(e, self.temb_w, self.temb_b, splitk=1)
        return proj.repeat_interleave(frames, dim=0).contiguous()

    def _res(self, r: _Resnet, x, temb_all, x2=None):
        b, f, hh, ww, c = x.shape
        y = r(x.view(b * f, hh, ww, c), temb_all, x2=None if x2 is None else x2.view(b * f, hh, ww, x2.shape[-1]))
        return y.view(b, f, hh, ww, -1)

    def forward(self, x: torch.Tensor, temb_all: torch.Tensor, kvs: List[torch.Tensor]) -> torch.Tensor:
        """x [B, F, H, W, 4] -> eps [B, F, H, W, 4]."""
        b, f, hh, ww, _ = x.shape
        it = iter(kvs)
        TemporalShardInfo.ranks = getattr(self, "frame_ranks", 1)
        h = ops.conv_direct(x.view(b * f, hh, ww, -1), self.conv_in_w, self.conv_in_b).view(b, f, hh, ww, -1)
        h = self.transformer_in(h)
        skips = [h]
        for blk in self.down:
            for j, r in enumerate(blk["res"]):
                h = blk["tc"][j](self._res(r, h, temb_all))
                if blk["att"] is not None:
                    h = blk["tatt"][j](blk["att"][j](h, next(it)))
                skips.append(h)
            if blk["down"] is not None:
                bb, ff, h2, w2, c = h.shape
                y = ops.conv3x3_strided(h.view(bb * ff, h2, w2, c), blk["down"][0], blk["down"][1], stride=2)
                h = y.view(bb, ff, y.shape[1], y.shape[2], -1)
                skips.append(h)
        m = self.mid
        h = m["tc"][0](self._res(m["res"][0], h, temb_all))
        h = m["tatt"](m["att"](h, next(it)))
        h = m["tc"][1](self._res(m["res"][1], h, temb_all))
        for blk in self.up:
            for j, r in enumerate(blk["res"]):
                h = blk["tc"][j](self._res(r, h, temb_all, x2=skips.pop()))
                if blk["att"] is not None:
                    h = blk["tatt"][j](blk["att"][j](h, next(it)))
            if blk["up"] is not None:
                bb, ff, h2, w2, c = h.shape
                y = ops.conv3x3(h.view(bb * ff, h2, w2, c), blk["up"][0], blk["up"][1], upsample=True)   # nearest 2x folded into the conv
                h = y.view(bb, ff, 2 * h2, 2 * w2, -1)
        bb, ff, h2, w2, c = h.shape
        y = ops.groupnorm(h.view(bb * ff, h2 * w2, c), self.no_w, self.no_b, self.cfg.norm_groups, self.cfg.norm_eps, "silu")
        y = ops.conv_direct(y.view(bb * ff, h2, w2, c), self.conv_out_w, self.conv_out_b)
        return y.view(bb, ff, h2, w2, -1)
