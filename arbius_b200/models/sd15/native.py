"""SD-1.5 on the sm_100a kernels: UNet, CLIP text encoder, VAE decoder.

The engines consume the SAME state dict as the eager references (diffusers / transformers key
layout, i.e. what upstream safetensors contain) and repack once at load:
  conv weights  [Cout, Cin, 3, 3] -> [Cout, 9*Cin] (kh, kw, cin)       (TMA implicit-GEMM order)
  q/k/v weights -> one fused [3C, C] (self) / [2C, ctx] (cross) matrix   (one GEMM, strided views)
  GEGLU proj    -> value/gate rows interleaved in blocks of 16           (gate fused in the epilogue)
  time_emb_proj of all resnets -> one [sum(Cout), 1280] matrix           (one GEMM per step)
Activations are channels-last end to end: [B, H, W, C] == [B, HW, C] tokens, so convs, 1x1
projections, attention and norms chain without a single permute or concat copy.
Cross-attention K/V depend only on the prompt and are computed once per task, not per step.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ... import ops
from ...ops import misc
from .clip_ref import ClipTextConfig
from .unet_ref import UNetConfig
from .vae_ref import VaeConfig

SD = Dict[str, torch.Tensor]


def _get(sd: SD, key: str, dtype, device) -> torch.Tensor:
    return sd[key].detach().to(device=device, dtype=dtype).contiguous()


class _Resnet:
    def __init__(self, sd: SD, p: str, dtype, device, eps: float, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.n1w, self.n1b = g("norm1.weight"), g("norm1.bias")
        self.c1w, self.c1b = ops.pack_conv_weight(g("conv1.weight")), g("conv1.bias")
        self.n2w, self.n2b = g("norm2.weight"), g("norm2.bias")
        self.c2w, self.c2b = ops.pack_conv_weight(g("conv2.weight")), g("conv2.bias")
        self.cout = self.c1w.shape[0]
        self.has_temb = (p + "time_emb_proj.weight") in sd
        self.tw = g("time_emb_proj.weight") if self.has_temb else None
        self.tb = g("time_emb_proj.bias") if self.has_temb else None
        self.temb_off = 0
        self.sw = None
        if (p + "conv_shortcut.weight") in sd:
            w = g("conv_shortcut.weight")
            self.sw, self.sb = w.reshape(w.shape[0], w.shape[1]).contiguous(), g("conv_shortcut.bias")
        self.eps, self.groups = eps, groups

    def __call__(self, x: torch.Tensor, temb_all: Optional[torch.Tensor], x2: Optional[torch.Tensor] = None):
        h = ops.groupnorm(x, self.n1w, self.n1b, self.groups, self.eps, "silu", x2=x2)
        rb = temb_all[:, self.temb_off:self.temb_off + self.cout] if (temb_all is not None and self.has_temb) else None
        h = ops.conv3x3(h, self.c1w, self.c1b, rowbias=rb)
        h = ops.groupnorm(h, self.n2w, self.n2b, self.groups, self.eps, "silu")
        if self.sw is not None:
            sc = ops.gemm(x, self.sw, self.sb, x2=x2)
        else:
            assert x2 is None
            sc = x
        return ops.conv3x3(h, self.c2w, self.c2b, residual=sc)


class _Transformer:
    def __init__(self, sd: SD, p: str, dtype, device, heads: int, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.heads, self.groups = heads, groups
        self.nw, self.nb = g("norm.weight"), g("norm.bias")
        w = g("proj_in.weight")
        self.c = w.shape[0]
        self.piw, self.pib = w.reshape(self.c, -1).contiguous(), g("proj_in.bias")
        w = g("proj_out.weight")
        self.pow, self.pob = w.reshape(self.c, -1).contiguous(), g("proj_out.bias")
        t = "transformer_blocks.0."
        self.ln1 = (g(t + "norm1.weight"), g(t + "norm1.bias"))
        self.ln2 = (g(t + "norm2.weight"), g(t + "norm2.bias"))
        self.ln3 = (g(t + "norm3.weight"), g(t + "norm3.bias"))
        self.wqkv = torch.cat([g(t + "attn1.to_q.weight"), g(t + "attn1.to_k.weight"), g(t + "attn1.to_v.weight")], 0).contiguous()
        self.wo1, self.bo1 = g(t + "attn1.to_out.0.weight"), g(t + "attn1.to_out.0.bias")
        self.wq2 = g(t + "attn2.to_q.weight")
        self.wkv2 = torch.cat([g(t + "attn2.to_k.weight"), g(t + "attn2.to_v.weight")], 0).contiguous()
        self.wo2, self.bo2 = g(t + "attn2.to_out.0.weight"), g(t + "attn2.to_out.0.bias")
        self.wff1, self.bff1 = ops.pack_geglu_weight(g(t + "ff.net.0.proj.weight"), g(t + "ff.net.0.proj.bias"))
        self.wff2, self.bff2 = g(t + "ff.net.2.weight"), g(t + "ff.net.2.bias")

    def context_kv(self, ctx: torch.Tensor) -> torch.Tensor:
        """[B, 77, ctx_dim] -> fused K|V projection [B, 77, 2C] (once per task)."""
        return ops.gemm(ctx, self.wkv2)

    def __call__(self, x: torch.Tensor, kv: torch.Tensor) -> torch.Tensor:
        b, hh, ww, c = x.shape
        l, h, d = hh * ww, self.heads, c // self.heads
        t = ops.groupnorm(x, self.nw, self.nb, self.groups, 1e-6, None)
        t = ops.gemm(t.view(b * l, c), self.piw, self.pib)                       # [B*L, C]
        # self-attention
        n = ops.layernorm(t, *self.ln1)
        qkv = ops.gemm(n, self.wqkv).view(b, l, 3 * c)
        q, k, v = (qkv[..., i * c:(i + 1) * c].view(b, l, h, d) for i in range(3))
        a = ops.flash_attn(q, k, v)
        t = ops.gemm(a.view(b * l, c), self.wo1, self.bo1, residual=t)
        # cross-attention (K/V precomputed)
        n = ops.layernorm(t, *self.ln2)
        q = ops.gemm(n, self.wq2).view(b, l, h, d)
        lk = kv.shape[1]
        k2, v2 = kv[..., :c].view(b, lk, h, d), kv[..., c:].view(b, lk, h, d)
        a = ops.flash_attn(q, k2, v2)
        t = ops.gemm(a.view(b * l, c), self.wo2, self.bo2, residual=t)
        # GEGLU feed-forward
        n = ops.layernorm(t, *self.ln3)
        f = ops.gemm(n, self.wff1, self.bff1, geglu=True)
        t = ops.gemm(f, self.wff2, self.bff2, residual=t)
        return ops.gemm(t, self.pow, self.pob, residual=x.view(b * l, c)).view(b, hh, ww, c)


class NativeUNet:
    def __init__(self, sd: SD, cfg: UNetConfig = UNetConfig(), dtype=torch.float16, device="cuda"):
        self.cfg, self.dtype, self.device = cfg, dtype, torch.device(device)
        g = lambda k: _get(sd, k, dtype, device)
        ch = cfg.block_out_channels
        self.conv_in_w = ops.DirectConvWeight(g("conv_in.weight").permute(0, 2, 3, 1))
        self.conv_in_b = g("conv_in.bias")
        self.te = (g("time_embedding.linear_1.weight"), g("time_embedding.linear_1.bias"),
                   g("time_embedding.linear_2.weight"), g("time_embedding.linear_2.bias"))
        mk_r = lambda p: _Resnet(sd, p, dtype, device, cfg.norm_eps, cfg.norm_groups)
        mk_t = lambda p: _Transformer(sd, p, dtype, device, cfg.attention_head_dim, cfg.norm_groups)
        self.down = []
        for i in range(len(ch)):
            blk = {"res": [mk_r(f"down_blocks.{i}.resnets.{j}.") for j in range(cfg.layers_per_block)],
                   "att": [mk_t(f"down_blocks.{i}.attentions.{j}.") for j in range(cfg.layers_per_block)]
                   if cfg.attention_levels[i] else None, "down": None}
            if i < len(ch) - 1:
                p = f"down_blocks.{i}.downsamplers.0.conv."
                blk["down"] = (ops.pack_conv_weight(g(p + "weight")), g(p + "bias"))
            self.down.append(blk)
        self.mid_res = [mk_r("mid_block.resnets.0."), mk_r("mid_block.resnets.1.")]
        self.mid_att = mk_t("mid_block.attentions.0.")
        self.up = []
        rev_attn = list(reversed(cfg.attention_levels))
        for i in range(len(ch)):
            n = cfg.layers_per_block + 1
            blk = {"res": [mk_r(f"up_blocks.{i}.resnets.{j}.") for j in range(n)],
                   "att": [mk_t(f"up_blocks.{i}.attentions.{j}.") for j in range(n)] if rev_attn[i] else None,
                   "up": None}
            if i < len(ch) - 1:
                p = f"up_blocks.{i}.upsamplers.0.conv."
                blk["up"] = (ops.pack_conv_weight(g(p + "weight")), g(p + "bias"))
            self.up.append(blk)
        self.no_w, self.no_b = g("conv_norm_out.weight"), g("conv_norm_out.bias")
        self.conv_out_w = ops.DirectConvWeight(g("conv_out.weight").permute(0, 2, 3, 1))
        self.conv_out_b = g("conv_out.bias")
        # one fused time-embedding projection for every resnet
        res_all: List[_Resnet] = []
        for blk in self.down:
            res_all += blk["res"]
        res_all += self.mid_res
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
        self.transformers: List[_Transformer] = []
        for blk in self.down:
            self.transformers += blk["att"] or []
        self.transformers.append(self.mid_att)
        for blk in self.up:
            self.transformers += blk["att"] or []

    def precompute_context(self, ctx: torch.Tensor) -> List[torch.Tensor]:
        return [t.context_kv(ctx) for t in self.transformers]

    def time_projections(self, t: torch.Tensor) -> torch.Tensor:
        """t [B] fp32 timesteps -> silu(time_embedding(t)) projected for every resnet: [B, sum(Cout)]."""
        e = ops.timestep_embedding(t, self.cfg.block_out_channels[0], self.dtype, self.cfg.flip_sin_to_cos,
                                   self.cfg.freq_shift)
        e = ops.gemm(e, self.te[0], self.te[1], act="silu", splitk=1)
        e = ops.gemm(e, self.te[2], self.te[3], act="silu", splitk=1)      # silu(temb): all consumers apply it
        return ops.gemm(e, self.temb_w, self.temb_b, splitk=1)

    def forward(self, x: torch.Tensor, temb_all: torch.Tensor, kvs: List[torch.Tensor]) -> torch.Tensor:
        """x [B, H, W, 4] -> eps [B, H, W, 4] (model dtype, channels-last)."""
        it = iter(kvs)
        h = ops.conv_direct(x, self.conv_in_w, self.conv_in_b)
        skips = [h]
        for blk in self.down:
            for j, r in enumerate(blk["res"]):
                h = r(h, temb_all)
                if blk["att"] is not None:
                    h = blk["att"][j](h, next(it))
                skips.append(h)
            if blk["down"] is not None:
                h = ops.conv3x3_strided(h, blk["down"][0], blk["down"][1], stride=2)
                skips.append(h)
        h = self.mid_res[0](h, temb_all)
        h = self.mid_att(h, next(it))
        h = self.mid_res[1](h, temb_all)
        for blk in self.up:
            for j, r in enumerate(blk["res"]):
                h = r(h, temb_all, x2=skips.pop())
                if blk["att"] is not None:
                    h = blk["att"][j](h, next(it))
            if blk["up"] is not None:
                h = ops.conv3x3(h, blk["up"][0], blk["up"][1], upsample=True)   # nearest 2x folded into the conv
        h = ops.groupnorm(h, self.no_w, self.no_b, self.cfg.norm_groups, self.cfg.norm_eps, "silu")
        return ops.conv_direct(h, self.conv_out_w, self.conv_out_b)


class NativeClipText:
    def __init__(self, sd: SD, cfg: ClipTextConfig = ClipTextConfig(), dtype=torch.float16, device="cuda"):
        self.cfg, self.dtype = cfg, dtype
        g = lambda k: _get(sd, "text_model." + k, dtype, device)
        self.tok, self.pos = g("embeddings.token_embedding.weight"), g("embeddings.position_embedding.weight")
        self.layers = []
        for i in range(cfg.layers):
            p = f"encoder.layers.{i}."
            self.layers.append(dict(
                ln1=(g(p + "layer_norm1.weight"), g(p + "layer_norm1.bias")),
                wqkv=torch.cat([g(p + f"self_attn.{n}_proj.weight") for n in "qkv"], 0).contiguous(),
                bqkv=torch.cat([g(p + f"self_attn.{n}_proj.bias") for n in "qkv"], 0).contiguous(),
                wo=g(p + "self_attn.out_proj.weight"), bo=g(p + "self_attn.out_proj.bias"),
                ln2=(g(p + "layer_norm2.weight"), g(p + "layer_norm2.bias")),
                w1=g(p + "mlp.fc1.weight"), b1=g(p + "mlp.fc1.bias"),
                w2=g(p + "mlp.fc2.weight"), b2=g(p + "mlp.fc2.bias")))
        self.lnf = (g("final_layer_norm.weight"), g("final_layer_norm.bias"))

    def forward(self, ids: torch.Tensor) -> torch.Tensor:
        b, l = ids.shape
        c, h = self.cfg.hidden, self.cfg.heads
        x = misc.embedding(ids, self.tok, self.pos).view(b * l, c)
        for L in self.layers:
            n = ops.layernorm(x, *L["ln1"], eps=self.cfg.eps)
            qkv = ops.gemm(n, L["wqkv"], L["bqkv"], splitk=1).view(b, l, 3 * c)
            q, k, v = (qkv[..., i * c:(i + 1) * c].view(b, l, h, c // h) for i in range(3))
            a = ops.flash_attn(q, k, v, causal=True)
            x = ops.gemm(a.view(b * l, c), L["wo"], L["bo"], residual=x, splitk=1)
            n = ops.layernorm(x, *L["ln2"], eps=self.cfg.eps)
            f = ops.gemm(n, L["w1"], L["b1"], act=self.cfg.act, splitk=1)
            x = ops.gemm(f, L["w2"], L["b2"], residual=x, splitk=1)
        return ops.layernorm(x, *self.lnf, eps=self.cfg.eps).view(b, l, c)


class NativeVaeDecoder:
    def __init__(self, sd: SD, cfg: VaeConfig = VaeConfig(), dtype=torch.float16, device="cuda"):
        self.cfg, self.dtype = cfg, dtype
        g = lambda k: _get(sd, k, dtype, device)
        lc = cfg.latent_channels
        # post_quant_conv (1x1, lc->lc) padded to 8 output channels so the direct-conv kernel can vectorise
        pq = torch.zeros(8, 1, 1, lc, dtype=dtype, device=device)
        pq[:lc] = (g("post_quant_conv.weight") / cfg.scaling_factor).to(dtype).permute(0, 2, 3, 1)
        self.pq_w = ops.DirectConvWeight(pq)
        self.pq_b = torch.zeros(8, dtype=dtype, device=device)
        self.pq_b[:lc] = g("post_quant_conv.bias")
        w = g("decoder.conv_in.weight").permute(0, 2, 3, 1)
        ci = torch.zeros(w.shape[0], 3, 3, 8, dtype=dtype, device=device)
        ci[..., :lc] = w
        self.ci_w, self.ci_b = ops.DirectConvWeight(ci), g("decoder.conv_in.bias")
        gr = cfg.norm_groups
        mk_r = lambda p: _Resnet(sd, p, dtype, device, 1e-6, gr)
        self.mid_res = [mk_r("decoder.mid_block.resnets.0."), mk_r("decoder.mid_block.resnets.1.")]
        a = "decoder.mid_block.attentions.0."
        self.att = dict(nw=g(a + "group_norm.weight"), nb=g(a + "group_norm.bias"),
                        wqkv=torch.cat([g(a + f"to_{n}.weight") for n in "qkv"], 0).contiguous(),
                        bqkv=torch.cat([g(a + f"to_{n}.bias") for n in "qkv"], 0).contiguous(),
                        wo=g(a + "to_out.0.weight"), bo=g(a + "to_out.0.bias"))
        self.up = []
        nb = len(cfg.block_out_channels)
        for i in range(nb):
            blk = {"res": [mk_r(f"decoder.up_blocks.{i}.resnets.{j}.") for j in range(cfg.layers_per_block + 1)],
                   "up": None}
            if i < nb - 1:
                p = f"decoder.up_blocks.{i}.upsamplers.0.conv."
                blk["up"] = (ops.pack_conv_weight(g(p + "weight")), g(p + "bias"))
            self.up.append(blk)
        self.no_w, self.no_b = g("decoder.conv_norm_out.weight"), g("decoder.conv_norm_out.bias")
        self.co_w = ops.DirectConvWeight(g("decoder.conv_out.weight").permute(0, 2, 3, 1))
        self.co_b = g("decoder.conv_out.bias")

    def _attention(self, x: torch.Tensor) -> torch.Tensor:
        b, hh, ww, c = x.shape
        l = hh * ww
        A = self.att
        n = ops.groupnorm(x, A["nw"], A["nb"], self.cfg.norm_groups, 1e-6, None)
        qkv = ops.gemm(n.view(b * l, c), A["wqkv"], A["bqkv"]).view(b, l, 3 * c)
        if c <= 192:
            q, k, v = (qkv[..., i * c:(i + 1) * c].view(b, l, 1, c) for i in range(3))
            o = ops.flash_attn(q, k, v).view(b * l, c)
        else:
            # single 512-wide head: S = QK^T and PV as two tcgen05 GEMMs around a row softmax
            outs = []
            for bi in range(b):
                q = qkv[bi, :, :c].contiguous()
                k = qkv[bi, :, c:2 * c].contiguous()
                vt = misc.transpose2d(qkv[bi, :, 2 * c:].contiguous())
                s = ops.gemm(q, k, None, out_scale=c ** -0.5, splitk=1)
                p = misc.softmax_rows(s, 1.0, out=s)
                outs.append(ops.gemm(p, vt, None, splitk=1))
            o = torch.cat(outs, 0) if b > 1 else outs[0]
        return ops.gemm(o, A["wo"], A["bo"], residual=x.view(b * l, c)).view(b, hh, ww, c)

    def forward(self, latents: torch.Tensor) -> torch.Tensor:
        """latents [B, h, w, 4] (model dtype, channels-last, still scaled) -> image [B, 8h, 8w, 3] in [-1, 1]."""
        z = ops.conv_direct(latents, self.pq_w, self.pq_b)
        h = ops.conv_direct(z, self.ci_w, self.ci_b)
        h = self.mid_res[0](h, None)
        h = self._attention(h)
        h = self.mid_res[1](h, None)
        for blk in self.up:
            for r in blk["res"]:
                h = r(h, None)
            if blk["up"] is not None:
                h = ops.conv3x3(h, blk["up"][0], blk["up"][1], upsample=True)   # nearest 2x folded into the conv
        h = ops.groupnorm(h, self.no_w, self.no_b, self.cfg.norm_groups, 1e-6, "silu")
        return ops.conv_direct(h, self.co_w, self.co_b)
