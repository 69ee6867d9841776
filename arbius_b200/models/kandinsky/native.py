"""Kandinsky 2.1 on the sm_100a kernels: XLM-R text encoder, diffusion prior, UNet, MoVQ decoder.

Same state dicts as the eager restatement in ``ref.py``; weights are repacked once. Notable mappings:
  * scale-shift ResBlock norm -> one GroupNorm kernel with per-(sample, channel) scale/shift operands that are
    column slices of ONE fused embedding projection for all ResBlocks of the UNet (one GEMM per step);
  * attention over [image-embedding tokens | text tokens | pixels]: the encoder K/V rows are projected once per
    task into the head of a per-layer [B, T+L, 3C] buffer, the per-step QKV GEMM writes the pixel rows behind
    them, and flash attention reads q/k/v as strided views — no concat kernels;
  * up/down ResBlocks resample with the nearest-2x / 2x2-average kernels on channels-last tensors;
  * MoVQ SpatialNorm = affine-free GroupNorm followed by a modulate kernel that reads the two latent-resolution
    1x1-conv maps with nearest-neighbour addressing (the upsampled maps are never materialised).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ... import ops
from ...ops import extra, misc
from ..sd15.native import _get
from .ref import KUNetConfig, MoVQConfig, PriorConfig, XlmrConfig

SD = Dict[str, torch.Tensor]


def _lin(sd, p, dtype, device):
    return _get(sd, p + ".weight", dtype, device), _get(sd, p + ".bias", dtype, device)


def _conv1(w: torch.Tensor) -> torch.Tensor:
    return w.reshape(w.shape[0], w.shape[1]).contiguous()


# ------------------------------------------------------------------------------------------------ text encoder
class NativeXlmr:
    def __init__(self, sd: SD, cfg: XlmrConfig, dtype, device):
        self.cfg = cfg
        g = lambda k: _get(sd, k, dtype, device)
        self.tok, self.pos = g("tok.weight"), g("pos.weight")
        self.ln_emb = (g("ln_emb.weight"), g("ln_emb.bias"))
        self.layers = []
        for i in range(cfg.layers):
            p = f"layers.{i}."
            self.layers.append(dict(
                wqkv=torch.cat([g(p + f"{n}.weight") for n in "qkv"], 0).contiguous(),
                bqkv=torch.cat([g(p + f"{n}.bias") for n in "qkv"], 0).contiguous(),
                o=_lin(sd, p + "o", dtype, device), ln1=_lin(sd, p + "ln1", dtype, device),
                fc1=_lin(sd, p + "fc1", dtype, device), fc2=_lin(sd, p + "fc2", dtype, device),
                ln2=_lin(sd, p + "ln2", dtype, device)))
        self.proj = _lin(sd, "proj", dtype, device)

    def forward(self, ids: torch.Tensor):
        b, l = ids.shape
        c, h = self.cfg.hidden, self.cfg.heads
        x = misc.embedding(ids, self.tok, self.pos).view(b * l, c)
        x = ops.layernorm(x, *self.ln_emb)
        for L in self.layers:
            qkv = ops.gemm(x, L["wqkv"], L["bqkv"], splitk=1).view(b, l, 3 * c)
            q, k, v = (qkv[..., i * c:(i + 1) * c].view(b, l, h, c // h) for i in range(3))
            a = ops.flash_attn(q, k, v)
            x = ops.layernorm(ops.gemm(a.view(b * l, c), *L["o"], residual=x, splitk=1), *L["ln1"])
            f = ops.gemm(x, *L["fc1"], act="gelu", splitk=1)
            x = ops.layernorm(ops.gemm(f, *L["fc2"], residual=x, splitk=1), *L["ln2"])
        tokens = x.view(b, l, c)
        pooled = extra.global_avgpool(tokens)                        # mean over tokens
        return tokens, ops.gemm(pooled, *self.proj, splitk=1)


# ------------------------------------------------------------------------------------------------ prior
class NativePrior:
    def __init__(self, sd: SD, cfg: PriorConfig, dtype, device):
        self.cfg, self.dtype = cfg, dtype
        g = lambda k: _get(sd, k, dtype, device)
        self.text_enc_proj = _lin(sd, "text_enc_proj", dtype, device)
        self.text_emb_proj = _lin(sd, "text_emb_proj", dtype, device)
        self.te0, self.te2 = _lin(sd, "time_embed.0", dtype, device), _lin(sd, "time_embed.2", dtype, device)
        self.x_proj = _lin(sd, "x_proj", dtype, device)
        self.query, self.pos = g("query"), g("pos")
        self.blocks = []
        for i in range(cfg.layers):
            p = f"blocks.{i}."
            self.blocks.append(dict(ln1=_lin(sd, p + "ln_1", dtype, device), qkv=_lin(sd, p + "c_qkv", dtype, device),
                                    proj=_lin(sd, p + "c_proj", dtype, device), ln2=_lin(sd, p + "ln_2", dtype, device),
                                    fc=_lin(sd, p + "c_fc", dtype, device), proj2=_lin(sd, p + "c_proj2", dtype, device)))
        self.ln_final = _lin(sd, "ln_final", dtype, device)
        self.out_proj = _lin(sd, "out_proj", dtype, device)

    def text_part(self, text_tokens: torch.Tensor, text_emb: torch.Tensor):
        """Step-independent rows of the sequence (computed once per task)."""
        b, l, _ = text_tokens.shape
        tt = ops.gemm(text_tokens.reshape(b * l, -1), *self.text_enc_proj, splitk=1).view(b, l, -1)
        te = ops.gemm(text_emb, *self.text_emb_proj, splitk=1)
        return tt, te

    def forward(self, x: torch.Tensor, t: torch.Tensor, tt: torch.Tensor, te: torch.Tensor) -> torch.Tensor:
        c = self.cfg
        b, w, h = x.shape[0], c.width, c.heads
        l = c.text_len + 4
        temb = ops.timestep_embedding(t, w, self.dtype, False, 0.0)
        temb = ops.gemm(ops.gemm(temb, *self.te0, act="silu", splitk=1), *self.te2, splitk=1)
        xp = ops.gemm(x, *self.x_proj, splitk=1)
        seq = torch.cat([tt, te[:, None], temb[:, None], xp[:, None], self.query.expand(b, -1, -1)], dim=1)
        hdn = ops.binary(seq.contiguous().view(-1), self.pos.expand(b, -1, -1).contiguous().view(-1), "add").view(b * l, w)
        for B in self.blocks:
            n = ops.layernorm(hdn, *B["ln1"])
            qkv = ops.gemm(n, *B["qkv"], splitk=1).view(b, l, 3 * w)
            q, k, v = (qkv[..., i * w:(i + 1) * w].view(b, l, h, w // h) for i in range(3))
            a = ops.flash_attn(q, k, v, causal=True)
            hdn = ops.gemm(a.view(b * l, w), *B["proj"], residual=hdn, splitk=1)
            n = ops.layernorm(hdn, *B["ln2"])
            hdn = ops.gemm(ops.gemm(n, *B["fc"], act="gelu", splitk=1), *B["proj2"], residual=hdn, splitk=1)
        last = hdn.view(b, l, w)[:, -1].contiguous()
        return ops.gemm(ops.layernorm(last, *self.ln_final), *self.out_proj, splitk=1)


# ------------------------------------------------------------------------------------------------ UNet
class _KRes:
    def __init__(self, sd: SD, p: str, dtype, device, groups: int, up: bool, down: bool):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.groups, self.up, self.down = groups, up, down
        self.n1 = (g("in_layers.0.weight"), g("in_layers.0.bias"))
        self.c1w, self.c1b = ops.pack_conv_weight(g("in_layers.2.weight")), g("in_layers.2.bias")
        self.ew, self.eb = g("emb_layers.1.weight"), g("emb_layers.1.bias")
        self.n2 = (g("out_layers.0.weight"), g("out_layers.0.bias"))
        self.c2w, self.c2b = ops.pack_conv_weight(g("out_layers.3.weight")), g("out_layers.3.bias")
        self.cout = self.c1w.shape[0]
        self.sw = None
        if (p + "skip_connection.weight") in sd:
            self.sw, self.sb = _conv1(g("skip_connection.weight")), g("skip_connection.bias")
        self.emb_off = 0

    def _rs(self, t):
        return ops.upsample_nearest2x(t) if self.up else extra.avgpool2x(t)

    def __call__(self, x: torch.Tensor, emb_all: torch.Tensor, x2: Optional[torch.Tensor] = None) -> torch.Tensor:
        h = ops.groupnorm(x, *self.n1, self.groups, 1e-5, "silu", x2=x2)
        if self.up or self.down:
            assert x2 is None
            h, x = self._rs(h), self._rs(x)
        h = ops.conv3x3(h, self.c1w, self.c1b)
        ss = emb_all[:, self.emb_off:self.emb_off + 2 * self.cout]
        h = ops.groupnorm(h, *self.n2, self.groups, 1e-5, "silu", scale=ss[:, :self.cout], shift=ss[:, self.cout:])
        sc = ops.gemm(x, self.sw, self.sb, x2=x2) if self.sw is not None else x
        return ops.conv3x3(h, self.c2w, self.c2b, residual=sc)


class _KAttn:
    def __init__(self, sd: SD, p: str, dtype, device, heads: int, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.heads, self.groups = heads, groups
        self.norm = (g("norm.weight"), g("norm.bias"))
        self.wqkv, self.bqkv = _conv1(g("qkv.weight")), g("qkv.bias")
        self.wekv, self.bekv = _conv1(g("encoder_kv.weight")), g("encoder_kv.bias")
        self.wo, self.bo = _conv1(g("proj_out.weight")), g("proj_out.bias")
        self.c = self.wo.shape[0]
        self.buf: Optional[torch.Tensor] = None

    def prepare(self, enc: torch.Tensor, l: int) -> None:
        """enc [B, T, mc]: project the encoder tokens once per task into rows [0, T) of the K/V columns."""
        b, t, _ = enc.shape
        c = self.c
        self.buf = torch.zeros((b, t + l, 3 * c), dtype=enc.dtype, device=enc.device)
        self.t = t
        for i in range(b):
            ops.gemm(enc[i], self.wekv, self.bekv, out=self.buf[i, :t, c:], splitk=1)

    off = 0                                    # first sample of the batch held here (guidance parallelism: the rank)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        b, hh, ww, c = x.shape
        l, h, t = hh * ww, self.heads, self.t
        n = ops.groupnorm(x, *self.norm, self.groups, 1e-5, None).view(b, l, c)
        buf = self.buf[self.off:self.off + b]
        for i in range(b):
            ops.gemm(n[i], self.wqkv, self.bqkv, out=buf[i, t:], splitk=1)
        q = buf[:, t:, :c].view(b, l, h, c // h)
        k = buf[:, :, c:2 * c].view(b, t + l, h, c // h)
        v = buf[:, :, 2 * c:].view(b, t + l, h, c // h)
        a = ops.flash_attn(q, k, v)
        return ops.gemm(a.view(b * l, c), self.wo, self.bo, residual=x.view(b * l, c)).view(b, hh, ww, c)


class NativeKandinskyUNet:
    def __init__(self, sd: SD, cfg: KUNetConfig, dtype=torch.float16, device="cuda"):
        self.cfg, self.dtype = cfg, dtype
        g = lambda k: _get(sd, k, dtype, device)
        mc, gr = cfg.model_channels, cfg.groups
        self.te0, self.te2 = _lin(sd, "time_embed.0", dtype, device), _lin(sd, "time_embed.2", dtype, device)
        self.ln_model_n = _lin(sd, "ln_model_n", dtype, device)
        self.to_model_dim_n = _lin(sd, "to_model_dim_n", dtype, device)
        self.proj_n, self.img_layer = _lin(sd, "proj_n", dtype, device), _lin(sd, "img_layer", dtype, device)
        self.clip_to_seq = _lin(sd, "clip_to_seq", dtype, device)
        self.conv_in = (ops.DirectConvWeight(g("input_blocks.0.0.weight").permute(0, 2, 3, 1)), g("input_blocks.0.0.bias"))
        self.res: List[_KRes] = []
        self.attn: List[_KAttn] = []

        def parse(prefix: str, n_layers_max: int = 3):
            layers = []
            for j in range(n_layers_max):
                p = f"{prefix}{j}."
                if (p + "in_layers.0.weight") in sd:
                    # up/down blocks are the ones whose in/out channel counts match AND that sit where the config says;
                    # the structure is re-derived from the config below, so flags are patched after parsing
                    r = _KRes(sd, p, dtype, device, gr, False, False)
                    self.res.append(r)
                    layers.append(r)
                elif (p + "qkv.weight") in sd:
                    c = sd[p + "proj_out.weight"].shape[0]
                    a = _KAttn(sd, p, dtype, device, c // cfg.head_dim, gr)
                    self.attn.append(a)
                    layers.append(a)
            return layers
        # walk the same construction order as ref.KandinskyUNet to recover which blocks resample
        self.input_blocks, idx, ds = [], 1, 1
        for level, mult in enumerate(cfg.channel_mult):
            for _ in range(cfg.num_res_blocks):
                self.input_blocks.append(parse(f"input_blocks.{idx}."))
                idx += 1
            if level != len(cfg.channel_mult) - 1:
                blk = parse(f"input_blocks.{idx}.")
                blk[0].down = True
                self.input_blocks.append(blk)
                idx += 1
                ds *= 2
        self.middle = parse("middle_block.")
        self.output_blocks, idx = [], 0
        for level, mult in list(enumerate(cfg.channel_mult))[::-1]:
            for i in range(cfg.num_res_blocks + 1):
                blk = parse(f"output_blocks.{idx}.")
                if level and i == cfg.num_res_blocks:
                    blk[-1].up = True
                self.output_blocks.append(blk)
                idx += 1
        self.out_norm = (g("out.0.weight"), g("out.0.bias"))
        self.out_conv = (ops.DirectConvWeight(g("out.2.weight").permute(0, 2, 3, 1)), g("out.2.bias"))
        off = 0
        for r in self.res:
            r.emb_off = off
            off += 2 * r.cout
        self.emb_w = torch.cat([r.ew for r in self.res], 0).contiguous()
        self.emb_b = torch.cat([r.eb for r in self.res], 0).contiguous()
        for r in self.res:
            r.ew = r.eb = None

    def prepare(self, text_tokens, text_emb, image_emb, latent_hw: int) -> torch.Tensor:
        """Per-task conditioning: encoder K/V rows for every attention layer; returns the step-independent part of
        the embedding (proj_n(text) + img_layer(image))."""
        c = self.cfg
        b, l, _ = text_tokens.shape
        seq = ops.gemm(ops.layernorm(text_tokens.reshape(b * l, -1), *self.ln_model_n), *self.to_model_dim_n, splitk=1)
        img = ops.gemm(image_emb, *self.clip_to_seq, splitk=1).view(b, c.image_tokens, c.model_channels)
        enc = torch.cat([img, seq.view(b, l, -1)], dim=1).contiguous()
        # pixel counts per attention layer follow the UNet's resolution schedule
        sizes = self._attn_sizes(latent_hw)
        for a, n_pix in zip(self.attn, sizes):
            a.prepare(enc, n_pix)
        base = ops.gemm(text_emb, *self.proj_n, splitk=1)
        return ops.gemm(image_emb, *self.img_layer, residual=base, splitk=1)

    def _attn_sizes(self, latent_hw: int) -> List[int]:
        c, out, ds = self.cfg, [], 1
        for level, _ in enumerate(c.channel_mult):
            for _ in range(c.num_res_blocks):
                if ds in c.attention_ds:
                    out.append(latent_hw // (ds * ds))
            if level != len(c.channel_mult) - 1:
                ds *= 2
        out.append(latent_hw // (ds * ds))            # middle
        for level, _ in list(enumerate(c.channel_mult))[::-1]:
            for i in range(c.num_res_blocks + 1):
                if ds in c.attention_ds:
                    out.append(latent_hw // (ds * ds))
                if level and i == c.num_res_blocks:
                    ds //= 2
        return out

    def embeddings(self, t: torch.Tensor, cond_emb: torch.Tensor) -> torch.Tensor:
        """t [B] -> fused scale/shift projections for every ResBlock: [B, sum(2*Cout)]."""
        e = ops.timestep_embedding(t, self.cfg.model_channels, self.dtype, False, 0.0)
        e = ops.gemm(e, *self.te0, act="silu", splitk=1)
        e = ops.gemm(e, *self.te2, residual=cond_emb, splitk=1)
        e = ops.binary(e.view(-1), None, "add", act="silu").view(e.shape)
        return ops.gemm(e, self.emb_w, self.emb_b, splitk=1)

    @staticmethod
    def _run(layers, h, emb_all, x2=None):
        for m in layers:
            if isinstance(m, _KRes):
                h = m(h, emb_all, x2=x2)
                x2 = None
            else:
                h = m(h)
        return h

    def forward(self, x: torch.Tensor, emb_all: torch.Tensor, sample_offset: int = 0) -> torch.Tensor:
        """x [B, H, W, 4] -> [B, H, W, 8]. ``sample_offset``: x holds samples [offset, offset+B) of the prepared batch."""
        for a in self.attn:
            a.off = sample_offset
        h = ops.conv_direct(x, *self.conv_in)
        hs = [h]
        for blk in self.input_blocks:
            h = self._run(blk, h, emb_all)
            hs.append(h)
        h = self._run(self.middle, h, emb_all)
        for blk in self.output_blocks:
            h = self._run(blk, h, emb_all, x2=hs.pop())
        h = ops.groupnorm(h, *self.out_norm, self.cfg.groups, 1e-5, "silu")
        return ops.conv_direct(h, *self.out_conv)


# ------------------------------------------------------------------------------------------------ MoVQ
class _SpatialNorm:
    def __init__(self, sd: SD, p: str, dtype, device, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.groups = groups
        self.nw, self.nb = g("norm_layer.weight"), g("norm_layer.bias")
        self.wy = ops.DirectConvWeight(g("conv_y.weight").permute(0, 2, 3, 1))
        self.by = g("conv_y.bias")
        self.wb = ops.DirectConvWeight(g("conv_b.weight").permute(0, 2, 3, 1))
        self.bb = g("conv_b.bias")

    def __call__(self, f: torch.Tensor, zq: torch.Tensor, act: Optional[str]) -> torch.Tensor:
        n = ops.groupnorm(f, self.nw, self.nb, self.groups, 1e-6, None)
        ymap = ops.conv_direct(zq, self.wy, self.by)           # latent resolution
        bmap = ops.conv_direct(zq, self.wb, self.bb)
        return extra.spatial_modulate(n, ymap, bmap, act)


class _MoVQRes:
    def __init__(self, sd: SD, p: str, dtype, device, groups: int):
        g = lambda k: _get(sd, p + k, dtype, device)
        self.n1 = _SpatialNorm(sd, p + "norm1.", dtype, device, groups)
        self.n2 = _SpatialNorm(sd, p + "norm2.", dtype, device, groups)
        self.c1 = (ops.pack_conv_weight(g("conv1.weight")), g("conv1.bias"))
        self.c2 = (ops.pack_conv_weight(g("conv2.weight")), g("conv2.bias"))
        self.sc = (_conv1(g("nin_shortcut.weight")), g("nin_shortcut.bias")) if (p + "nin_shortcut.weight") in sd else None

    def __call__(self, x, zq):
        h = ops.conv3x3(self.n1(x, zq, "silu"), *self.c1)
        h = self.n2(h, zq, "silu")
        sc = ops.gemm(x, *self.sc) if self.sc is not None else x
        return ops.conv3x3(h, *self.c2, residual=sc)


class NativeMoVQDecoder:
    def __init__(self, sd: SD, cfg: MoVQConfig, dtype=torch.float16, device="cuda"):
        self.cfg = cfg
        g = lambda k: _get(sd, k, dtype, device)
        zc, gr = cfg.z_channels, cfg.groups
        pq = torch.zeros(8, 1, 1, zc, dtype=dtype, device=device)
        pq[:zc] = (g("post_quant_conv.weight") / cfg.scaling).to(dtype).permute(0, 2, 3, 1)
        self.pq = (ops.DirectConvWeight(pq), torch.cat([g("post_quant_conv.bias"), torch.zeros(8 - zc, dtype=dtype, device=device)]))
        self.zc = zc

        def pad_in(w):   # [Cout, zc, k, k] -> [Cout, k, k, 8] (latent carried with 8 channels, 4 of them zero)
            w = w.permute(0, 2, 3, 1)
            out = torch.zeros(w.shape[:3] + (8,), dtype=dtype, device=device)
            out[..., :zc] = w
            return ops.DirectConvWeight(out)
        self._pad_in = pad_in
        self.conv_in = (pad_in(g("conv_in.weight")), g("conv_in.bias"))
        mk = lambda p: _MoVQRes(sd, p, dtype, device, gr)
        for r in ():
            pass
        self.mid1, self.mid2 = mk("mid_block_1."), mk("mid_block_2.")
        a = "mid_attn_1."
        self.attn_norm = _SpatialNorm(sd, a + "norm.", dtype, device, gr)
        self.attn = dict(wqkv=torch.cat([_conv1(g(a + f"{n}.weight")) for n in "qkv"], 0).contiguous(),
                         bqkv=torch.cat([g(a + f"{n}.bias") for n in "qkv"], 0).contiguous(),
                         wo=_conv1(g(a + "proj_out.weight")), bo=g(a + "proj_out.bias"))
        self.up = []
        for s in range(len(cfg.ch_mult)):
            blocks = [mk(f"up.{s}.block.{j}.") for j in range(cfg.num_res_blocks + 1)]
            upc = None
            if (f"up.{s}.upsample.weight") in sd:
                upc = (ops.pack_conv_weight(g(f"up.{s}.upsample.weight")), g(f"up.{s}.upsample.bias"))
            self.up.append((blocks, upc))
        self.norm_out = _SpatialNorm(sd, "norm_out.", dtype, device, gr)
        self.conv_out = (ops.DirectConvWeight(g("conv_out.weight").permute(0, 2, 3, 1)), g("conv_out.bias"))
        # SpatialNorm 1x1 convs read the 8-channel latent as well
        for sn in self._all_spatial_norms():
            for attr in ("wy", "wb"):
                w = getattr(sn, attr).t                           # [1, 1, zc, C] (few_in layout: kh, kw, cin, cout)
                w8 = torch.zeros((1, 1, 8, w.shape[-1]), dtype=dtype, device=device)
                w8[:, :, :zc] = w
                dw = getattr(sn, attr)
                dw.t, dw.cin = w8.contiguous(), 8

    def _all_spatial_norms(self):
        out = [self.attn_norm, self.norm_out]
        for r in [self.mid1, self.mid2] + [b for blocks, _ in self.up for b in blocks]:
            out += [r.n1, r.n2]
        return out

    def _attention(self, x, zq):
        b, hh, ww, c = x.shape
        l = hh * ww
        n = self.attn_norm(x, zq, None)
        A = self.attn
        qkv = ops.gemm(n.view(b * l, c), A["wqkv"], A["bqkv"]).view(b, l, 3 * c)
        if c <= 64:
            q, k, v = (qkv[..., i * c:(i + 1) * c].view(b, l, 1, c) for i in range(3))
            o = ops.flash_attn(q, k, v).view(b * l, c)
        else:
            outs = []
            for bi in range(b):
                q = qkv[bi, :, :c].contiguous()
                k = qkv[bi, :, c:2 * c].contiguous()
                vt = misc.transpose2d(qkv[bi, :, 2 * c:].contiguous())
                s = ops.gemm(q, k, None, out_scale=c ** -0.5, splitk=1)
                outs.append(ops.gemm(misc.softmax_rows(s, 1.0, out=s), vt, None, splitk=1))
            o = torch.cat(outs, 0) if b > 1 else outs[0]
        return ops.gemm(o, A["wo"], A["bo"], residual=x.view(b * l, c)).view(b, hh, ww, c)

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        """z [B, h, w, 4] -> image [B, H, W, 3] in [-1, 1]."""
        zq = ops.conv_direct(z, *self.pq)                         # [B, h, w, 8]
        h = ops.conv_direct(zq, *self.conv_in)
        h = self.mid1(h, zq)
        h = self._attention(h, zq)
        h = self.mid2(h, zq)
        for blocks, upc in self.up:
            for blk in blocks:
                h = blk(h, zq)
            if upc is not None:
                h = ops.conv3x3(ops.upsample_nearest2x(h), *upc)
        h = self.norm_out(h, zq, "silu")
        return ops.conv_direct(h, *self.conv_out)
