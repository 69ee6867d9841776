"""CLIP ViT-L/14 text encoder (12 layers, d=768, 12 heads, 77 tokens, causal, quick-GELU) — eager oracle.
Parameter names follow transformers' ``CLIPTextModel`` (``text_model.*``) so upstream checkpoints load
unchanged. The same class serves OpenCLIP ViT-H (SD-2 style: d=1024, 23 used layers, GELU) for
ZeroScope/DAMO by changing the config."""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class ClipTextConfig:
    vocab_size: int = 49408
    max_positions: int = 77
    hidden: int = 768
    layers: int = 12
    heads: int = 12
    intermediate: int = 3072
    act: str = "quick_gelu"
    eps: float = 1e-5

    @staticmethod
    def tiny() -> "ClipTextConfig":
        return ClipTextConfig(vocab_size=1000, hidden=64, layers=2, heads=2, intermediate=128)

    @staticmethod
    def openclip_h() -> "ClipTextConfig":
        return ClipTextConfig(hidden=1024, layers=23, heads=16, intermediate=4096, act="gelu")


def _act(x, kind):
    return x * torch.sigmoid(1.702 * x) if kind == "quick_gelu" else F.gelu(x)


class ClipAttention(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.heads = c.heads
        self.q_proj = nn.Linear(c.hidden, c.hidden)
        self.k_proj = nn.Linear(c.hidden, c.hidden)
        self.v_proj = nn.Linear(c.hidden, c.hidden)
        self.out_proj = nn.Linear(c.hidden, c.hidden)

    def forward(self, x):
        b, l, d = x.shape
        sh = lambda t: t.view(b, l, self.heads, d // self.heads).transpose(1, 2)
        o = F.scaled_dot_product_attention(sh(self.q_proj(x)), sh(self.k_proj(x)), sh(self.v_proj(x)), is_causal=True)
        return self.out_proj(o.transpose(1, 2).reshape(b, l, d))


class ClipMLP(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.fc1 = nn.Linear(c.hidden, c.intermediate)
        self.fc2 = nn.Linear(c.intermediate, c.hidden)
        self.act = c.act

    def forward(self, x):
        return self.fc2(_act(self.fc1(x), self.act))


class ClipLayer(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.self_attn = ClipAttention(c)
        self.layer_norm1 = nn.LayerNorm(c.hidden, eps=c.eps)
        self.mlp = ClipMLP(c)
        self.layer_norm2 = nn.LayerNorm(c.hidden, eps=c.eps)

    def forward(self, x):
        x = x + self.self_attn(self.layer_norm1(x))
        return x + self.mlp(self.layer_norm2(x))


class ClipEmbeddings(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.token_embedding = nn.Embedding(c.vocab_size, c.hidden)
        self.position_embedding = nn.Embedding(c.max_positions, c.hidden)

    def forward(self, ids):
        pos = torch.arange(ids.shape[1], device=ids.device)
        return self.token_embedding(ids) + self.position_embedding(pos)[None]


class ClipEncoder(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.layers = nn.ModuleList([ClipLayer(c) for _ in range(c.layers)])


class ClipTextTransformer(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.embeddings = ClipEmbeddings(c)
        self.encoder = ClipEncoder(c)
        self.final_layer_norm = nn.LayerNorm(c.hidden, eps=c.eps)


class CLIPTextModel(nn.Module):
    def __init__(self, cfg: ClipTextConfig = ClipTextConfig()):
        super().__init__()
        self.cfg = cfg
        self.text_model = ClipTextTransformer(cfg)

    def forward(self, ids: torch.Tensor) -> torch.Tensor:
        """ids [B, 77] int64 -> last hidden state [B, 77, hidden]."""
        tm = self.text_model
        x = tm.embeddings(ids)
        for layer in tm.encoder.layers:
            x = layer(x)
        return tm.final_layer_norm(x)
