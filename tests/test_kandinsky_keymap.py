"""Kandinsky 2.1 upstream checkpoint layouts -> this repository's modules (models/kandinsky/keymap.py) and the XLM-R
tokenizer. The upstream files are not reachable offline, so the state dicts here are synthetic but carry the upstream
NAMES, and the attention layouts are checked against small restatements of the upstream code (GLIDE-style
QKVMultiheadAttention / guided-diffusion AttentionBlock with encoder_kv: fused projection viewed as [heads, 3, d])."""
import math

import pytest
import torch
import torch.nn.functional as F

from arbius_b200.models.kandinsky import keymap
from arbius_b200.models.kandinsky.ref import (DiffusionPrior, KAttentionBlock, KandinskyUNet, KUNetConfig, MoVQConfig,
                                              MoVQDecoder, PriorBlock, PriorConfig, TextEncoderXLMR, XlmrConfig)


def _rand_like(sd, seed=0):
    g = torch.Generator().manual_seed(seed)
    return {k: torch.randn(v.shape, generator=g) * 0.1 for k, v in sd.items()}


def test_prior_block_matches_upstream_interleaved_attention():
    c = PriorConfig.tiny()
    w, h = c.width, c.heads
    up = {"c_qkv.weight": torch.randn(3 * w, w) * 0.2, "c_qkv.bias": torch.randn(3 * w) * 0.1,
          "c_proj.weight": torch.randn(w, w) * 0.2, "c_proj.bias": torch.randn(w) * 0.1}
    x = torch.randn(2, 9, w)

    def upstream_attention(x):                       # xf.py QKVMultiheadAttention (+ causal mask, as the prior uses it)
        b, n, _ = x.shape
        qkv = F.linear(x, up["c_qkv.weight"], up["c_qkv.bias"]).view(b, n, h, -1)
        q, k, v = torch.split(qkv, w // h, dim=-1)
        scale = 1 / math.sqrt(math.sqrt(w // h))
        wgt = torch.einsum("bthc,bshc->bhts", q * scale, k * scale)
        wgt = wgt + torch.full((n, n), float("-inf")).triu(1)
        a = torch.einsum("bhts,bshc->bthc", torch.softmax(wgt, -1), v).reshape(b, n, -1)
        return F.linear(a, up["c_proj.weight"], up["c_proj.bias"])
    blk = PriorBlock(c)
    sd = {f"model.transformer.resblocks.0.attn.{k}": v for k, v in up.items()}
    mapped = keymap.map_prior({**sd, **_prior_rest(c)}, h)
    blk.c_qkv.weight.data, blk.c_qkv.bias.data = mapped["blocks.0.c_qkv.weight"], mapped["blocks.0.c_qkv.bias"]
    blk.c_proj.weight.data, blk.c_proj.bias.data = mapped["blocks.0.c_proj.weight"], mapped["blocks.0.c_proj.bias"]
    q, k, v = blk.c_qkv(x).chunk(3, dim=-1)
    sh = lambda t: t.view(2, 9, h, w // h).transpose(1, 2)
    ours = blk.c_proj(F.scaled_dot_product_attention(sh(q), sh(k), sh(v), is_causal=True).transpose(1, 2).reshape(2, 9, w))
    assert torch.allclose(ours, upstream_attention(x), atol=1e-5)


def _prior_rest(c, text_dim=64):
    w = c.width
    names = {"text_enc_proj": (w, text_dim), "text_emb_proj": (w, c.clip_dim), "time_embed.0": (w, w), "time_embed.2": (w, w),
             "clip_img_proj": (w, c.clip_dim), "final_ln": (w,), "out_proj": (c.clip_dim, w)}
    sd = {}
    for n, shp in names.items():
        sd[f"model.{n}.weight"] = torch.randn(*shp) * 0.1
        sd[f"model.{n}.bias"] = torch.randn(shp[0]) * 0.1
    sd["model.prd_emb"] = torch.randn(1, 1, w)
    sd["model.positional_embedding"] = torch.randn(1, c.text_len + 4, w)
    return sd


def test_prior_upstream_names_load_strict():
    c = PriorConfig.tiny()
    sd = _prior_rest(c)
    for i in range(c.layers):
        p = f"model.transformer.resblocks.{i}."
        for n, shp in (("ln_1", (c.width,)), ("attn.c_qkv", (3 * c.width, c.width)), ("attn.c_proj", (c.width, c.width)),
                       ("ln_2", (c.width,)), ("mlp.c_fc", (4 * c.width, c.width)), ("mlp.c_proj", (c.width, 4 * c.width))):
            sd[p + n + ".weight"] = torch.randn(*shp) * 0.1
            sd[p + n + ".bias"] = torch.randn(shp[0]) * 0.1
    assert keymap.is_upstream_prior(sd)
    DiffusionPrior(c, text_dim=64).load_state_dict(keymap.map_prior(sd, c.heads), strict=True)


def test_unet_attention_matches_upstream_interleaved_layout_with_encoder_kv():
    c, hd, enc_c, groups = 64, 32, 48, 8
    heads = c // hd
    blk = KAttentionBlock(c, hd, enc_c, groups)
    up = _rand_like({k: v for k, v in blk.state_dict().items()}, 3)
    x, enc = torch.randn(2, c, 4, 4), torch.randn(2, enc_c, 7)

    def upstream(x, enc):                            # guided-diffusion AttentionBlock + QKVAttention(encoder_kv)
        b, ch, hh, ww = x.shape
        xf = x.reshape(b, ch, -1)
        qkv = F.conv1d(F.group_norm(xf, groups, up["norm.weight"], up["norm.bias"]), up["qkv.weight"], up["qkv.bias"])
        ekv = F.conv1d(enc, up["encoder_kv.weight"], up["encoder_kv.bias"])
        d = ch // heads
        q, k, v = qkv.reshape(b * heads, d * 3, -1).split(d, dim=1)
        ek, ev = ekv.reshape(b * heads, d * 2, -1).split(d, dim=1)
        k, v = torch.cat([ek, k], dim=-1), torch.cat([ev, v], dim=-1)
        scale = 1 / math.sqrt(math.sqrt(d))
        wgt = torch.softmax(torch.einsum("bct,bcs->bts", q * scale, k * scale), dim=-1)
        a = torch.einsum("bts,bcs->bct", wgt, v).reshape(b, -1, xf.shape[-1])
        return (xf + F.conv1d(a, up["proj_out.weight"], up["proj_out.bias"])).reshape(b, ch, hh, ww)
    blk.load_state_dict(keymap.map_unet({f"input_blocks.4.1.{k}": v for k, v in up.items()}, hd) and
                        {k: v for k, v in ((kk.split("input_blocks.4.1.")[1], vv) for kk, vv in
                                           keymap.map_unet({f"input_blocks.4.1.{k}": v for k, v in up.items()}, hd).items())})
    assert torch.allclose(blk(x, enc), upstream(x, enc), atol=1e-5)


def test_unet_names_are_upstream_names():
    """The decoder file needs no renaming: every parameter of the restated UNet has a guided-diffusion name."""
    sd = KandinskyUNet(KUNetConfig.tiny()).state_dict()
    assert "input_blocks.0.0.weight" in sd and "time_embed.0.weight" in sd and "out.2.weight" in sd
    assert any(k.endswith("encoder_kv.weight") for k in sd) and any(k.endswith("in_layers.2.weight") for k in sd)
    assert set(keymap.map_unet(sd, 32)) == set(sd)


def test_text_encoder_hf_names_load_strict_and_fold_position_offset():
    c = XlmrConfig.tiny()
    hid, inter = c.hidden, c.intermediate
    sd = {"model.transformer.embeddings.word_embeddings.weight": torch.randn(c.vocab_size, hid),
          "model.transformer.embeddings.position_embeddings.weight": torch.randn(c.max_positions + 2, hid),
          "model.transformer.embeddings.token_type_embeddings.weight": torch.randn(1, hid),
          "model.transformer.embeddings.LayerNorm.weight": torch.randn(hid), "model.transformer.embeddings.LayerNorm.bias": torch.randn(hid),
          "model.transformer.pooler.dense.weight": torch.randn(hid, hid), "model.transformer.pooler.dense.bias": torch.randn(hid),
          "model.LinearTransformation.weight": torch.randn(c.out_dim, hid), "model.LinearTransformation.bias": torch.randn(c.out_dim)}
    for i in range(c.layers):
        p = f"model.transformer.encoder.layer.{i}."
        for n, shp in (("attention.self.query", (hid, hid)), ("attention.self.key", (hid, hid)), ("attention.self.value", (hid, hid)),
                       ("attention.output.dense", (hid, hid)), ("attention.output.LayerNorm", (hid,)),
                       ("intermediate.dense", (inter, hid)), ("output.dense", (hid, inter)), ("output.LayerNorm", (hid,))):
            sd[p + n + ".weight"] = torch.randn(*shp) * 0.1
            sd[p + n + ".bias"] = torch.randn(shp[0]) * 0.1
    assert keymap.is_upstream_text_encoder(sd)
    mapped = keymap.map_text_encoder(sd, c.max_positions)
    TextEncoderXLMR(c).load_state_dict(mapped, strict=True)
    pos, tt = sd["model.transformer.embeddings.position_embeddings.weight"], sd["model.transformer.embeddings.token_type_embeddings.weight"]
    assert torch.equal(mapped["pos.weight"][0], pos[2] + tt[0]) and torch.equal(mapped["pos.weight"][5], pos[7] + tt[0])


def test_movq_decoder_names_and_level_order():
    c = MoVQConfig.tiny()
    ours = MoVQDecoder(c).state_dict()
    L = len(c.ch_mult)
    up = {}
    for k, v in ours.items():                        # invert the map: build a taming-style dict from ours
        if k.startswith("post_quant_conv."):
            up[k] = v
            continue
        r = k.replace("mid_block_1.", "mid.block_1.").replace("mid_attn_1.", "mid.attn_1.").replace("mid_block_2.", "mid.block_2.")
        if r.startswith("up."):
            i, rest = r.split(".", 2)[1:]
            r = f"up.{L - 1 - int(i)}." + rest.replace("upsample.", "upsample.conv.")
        up["decoder." + r] = v
    up["encoder.conv_in.weight"] = torch.zeros(1)    # dropped
    up["quantize.embedding.weight"] = torch.zeros(1)
    assert keymap.is_upstream_movq(up)
    mapped = keymap.map_movq(up, L)
    assert set(mapped) == set(ours) and all(torch.equal(mapped[k], ours[k]) for k in ours)
    MoVQDecoder(c).load_state_dict(mapped, strict=True)


def test_xlmr_tokenizer_frames_like_fairseq(tmp_path):
    spm = pytest.importorskip("sentencepiece")
    corpus = tmp_path / "c.txt"
    corpus.write_text("\n".join(["arbius test cat", "a panda eating bamboo on a rock", "kandinsky two point one",
                                 "the quick brown fox jumps over the lazy dog"] * 20))
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=str(tmp_path / "sentencepiece.bpe"), vocab_size=36, hard_vocab_limit=False,
                                   model_type="unigram", unk_id=0, bos_id=1, eos_id=2, pad_id=-1)   # XLM-R: <unk> <s> </s>
    from arbius_b200.models.tokenizer import HashTokenizer, XlmrTokenizer, load_xlmr_tokenizer
    tok = load_xlmr_tokenizer(tmp_path, 250002)
    assert isinstance(tok, XlmrTokenizer)
    ids = tok.encode("arbius test cat")
    assert len(ids) == 77 and ids[0] == 0 and 2 in ids and ids[-1] == 1           # <s> ... </s> <pad>...
    body = ids[1:ids.index(2)]
    sp = spm.SentencePieceProcessor(model_file=str(tmp_path / "sentencepiece.bpe.model"))
    assert body == [i + 1 for i in sp.encode("arbius test cat")]                   # fairseq offset
    assert tok.encode("") == [0, 2] + [1] * 75
    assert isinstance(load_xlmr_tokenizer(None, 1000), HashTokenizer)
