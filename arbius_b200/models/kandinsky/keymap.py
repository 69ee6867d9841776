"""Upstream checkpoint layouts of Kandinsky 2.1 -> the module layouts of ``ref.py`` (and, through them, the native
engines). The reference names the model only by container digest and git URL (templates/kandinsky2.json ``meta.git``:
``kasumi-1/Kandinsky-2@aa5ee2f``); the four files that repository loads are

  text_encoder.ckpt   MultilingualCLIP: HF ``XLMRobertaModel`` under ``model.transformer.*`` plus
                      ``model.LinearTransformation`` (pooled projection)
  prior_fp16.ckpt     diffusion prior: GLIDE-style ``xf.py`` transformer under ``model.transformer.resblocks.N.*``
  decoder_fp16.ckpt   guided-diffusion UNet (``input_blocks / middle_block / output_blocks``) with ``encoder_kv``
  movq_final.ckpt     MoVQ autoencoder; only ``decoder.*`` and ``post_quant_conv`` are needed for text-to-image

[NOT IN TREE] None of these files is reachable offline, so the key names and the two layout facts below come from the
upstream source as published, not from the reference tree; ``tests/test_kandinsky_keymap.py`` checks the maps against
synthetic state dicts that carry the upstream names and against small restatements of the upstream attention code.

Layout facts that are more than a rename:
  * GLIDE / guided-diffusion attention (``QKVAttentionLegacy``-style, used by both the UNet's ``AttentionBlock`` and the
    prior's ``QKVMultiheadAttention``) views the fused projection as [heads, 3, head_dim]: the rows of ``qkv`` /
    ``c_qkv`` are interleaved per head (q_h, k_h, v_h), and ``encoder_kv`` as [heads, 2, head_dim]. The engines here
    use [3, heads, head_dim] (Q block, K block, V block) so that K and V are column ranges one TMA tensor map can walk;
    ``_deinterleave`` permutes the rows once at load. The checkpoint on disk is untouched.
  * XLM-R adds ``token_type_embeddings[0]`` to every position and numbers positions from ``padding_idx + 1 = 2``; both
    are folded into the position table at load (row i of ``pos`` = position_embeddings[i + 2] + token_type[0]).
"""
from __future__ import annotations

import re
from typing import Dict

import torch

SD = Dict[str, torch.Tensor]


def _deinterleave(w: torch.Tensor, heads: int, parts: int) -> torch.Tensor:
    """Rows [heads, parts, d] -> [parts, heads, d] (weight [rows, ...] or bias [rows])."""
    rows = w.shape[0]
    d = rows // (heads * parts)
    assert heads * parts * d == rows, (rows, heads, parts)
    return w.reshape(heads, parts, d, *w.shape[1:]).transpose(0, 1).reshape(rows, *w.shape[1:]).contiguous()


def is_upstream_text_encoder(sd: SD) -> bool:
    return any(k.startswith("model.transformer.embeddings.") for k in sd)


def map_text_encoder(sd: SD, max_positions: int) -> SD:
    """MultilingualCLIP (HF XLM-R) -> ``TextEncoderXLMR``."""
    p = "model.transformer."
    out: SD = {"tok.weight": sd[p + "embeddings.word_embeddings.weight"]}
    pos = sd[p + "embeddings.position_embeddings.weight"]
    tt = sd[p + "embeddings.token_type_embeddings.weight"][0]
    table = pos[2:] + tt                                   # positions start at padding_idx + 1 = 2
    if table.shape[0] < max_positions:
        table = torch.cat([table, table.new_zeros(max_positions - table.shape[0], table.shape[1])])
    out["pos.weight"] = table[:max_positions].contiguous()
    for s in ("weight", "bias"):
        out[f"ln_emb.{s}"] = sd[p + f"embeddings.LayerNorm.{s}"]
        out[f"proj.{s}"] = sd[f"model.LinearTransformation.{s}"]
    layer = re.compile(r"model\.transformer\.encoder\.layer\.(\d+)\.(.+)\.(weight|bias)$")
    names = {"attention.self.query": "q", "attention.self.key": "k", "attention.self.value": "v",
             "attention.output.dense": "o", "attention.output.LayerNorm": "ln1", "intermediate.dense": "fc1",
             "output.dense": "fc2", "output.LayerNorm": "ln2"}
    for k, v in sd.items():
        m = layer.match(k)
        if m and m.group(2) in names:
            out[f"layers.{m.group(1)}.{names[m.group(2)]}.{m.group(3)}"] = v
    return out


def is_upstream_prior(sd: SD) -> bool:
    return any(".transformer.resblocks." in k for k in sd)


def map_prior(sd: SD, heads: int) -> SD:
    """xf.py prior (``model.*``) -> ``DiffusionPrior``."""
    pre = "model." if any(k.startswith("model.") for k in sd) else ""
    g = lambda k: sd[pre + k]
    out: SD = {}
    for s in ("weight", "bias"):
        out[f"text_enc_proj.{s}"] = g(f"text_enc_proj.{s}")
        out[f"text_emb_proj.{s}"] = g(f"text_emb_proj.{s}")
        out[f"time_embed.0.{s}"] = g(f"time_embed.0.{s}")
        out[f"time_embed.2.{s}"] = g(f"time_embed.2.{s}")
        out[f"x_proj.{s}"] = g(f"clip_img_proj.{s}")
        out[f"ln_final.{s}"] = g(f"final_ln.{s}")
        out[f"out_proj.{s}"] = g(f"out_proj.{s}")
    out["query"] = g("prd_emb").reshape(1, 1, -1)
    out["pos"] = g("positional_embedding").reshape(1, -1, out["query"].shape[-1])
    blk = re.compile(re.escape(pre) + r"transformer\.resblocks\.(\d+)\.(.+)\.(weight|bias)$")
    names = {"ln_1": "ln_1", "attn.c_qkv": "c_qkv", "attn.c_proj": "c_proj", "ln_2": "ln_2", "mlp.c_fc": "c_fc",
             "mlp.c_proj": "c_proj2"}
    for k, v in sd.items():
        m = blk.match(k)
        if m and m.group(2) in names:
            if m.group(2) == "attn.c_qkv":
                v = _deinterleave(v, heads, 3)
            out[f"blocks.{m.group(1)}.{names[m.group(2)]}.{m.group(3)}"] = v
    return out


def map_unet(sd: SD, head_dim: int) -> SD:
    """guided-diffusion UNet: names are kept; fused attention projections go from per-head interleaved to blockwise."""
    out: SD = {}
    for k, v in sd.items():
        if re.search(r"\.qkv\.(weight|bias)$", k):
            v = _deinterleave(v, v.shape[0] // (3 * head_dim), 3)
        elif re.search(r"\.encoder_kv\.(weight|bias)$", k):
            v = _deinterleave(v, v.shape[0] // (2 * head_dim), 2)
        out[k] = v
    return out


def is_upstream_movq(sd: SD) -> bool:
    return any(k.startswith("decoder.mid.") for k in sd)


def map_movq(sd: SD, levels: int) -> SD:
    """MoVQ (taming-style ``decoder.*``) -> ``MoVQDecoder``: ``up`` is indexed by resolution level upstream (0 = highest)
    and from the lowest resolution up here; encoder / quantiser / loss keys are dropped."""
    out: SD = {}
    for k, v in sd.items():
        if k.startswith("post_quant_conv."):
            out[k] = v
            continue
        if not k.startswith("decoder."):
            continue
        r = k[len("decoder."):]
        r = r.replace("mid.block_1.", "mid_block_1.").replace("mid.attn_1.", "mid_attn_1.").replace("mid.block_2.", "mid_block_2.")
        m = re.match(r"up\.(\d+)\.(.+)$", r)
        if m:
            r = f"up.{levels - 1 - int(m.group(1))}." + m.group(2).replace("upsample.conv.", "upsample.")
        out[r] = v
    return out
