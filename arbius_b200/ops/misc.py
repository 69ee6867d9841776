"""Small helpers: row softmax, 2-D transpose, embedding gather (csrc/norm.cu tail)."""
from __future__ import annotations

import torch

from . import native


def softmax_rows(x: torch.Tensor, scale: float = 1.0, out: torch.Tensor | None = None) -> torch.Tensor:
    lib = native.load()
    cols = x.shape[-1]
    rows = x.numel() // cols
    y = out if out is not None else torch.empty_like(x)
    native.check(lib.ab_softmax_rows(x.data_ptr(), y.data_ptr(), rows, cols, scale, native.dtype_code(x.dtype),
                                     native.stream_ptr()), "ab_softmax_rows")
    return y


def transpose2d(x: torch.Tensor) -> torch.Tensor:
    """[..., R, C] -> [..., C, R] contiguous."""
    lib = native.load()
    r, c = x.shape[-2], x.shape[-1]
    batch = x.numel() // (r * c)
    y = torch.empty(x.shape[:-2] + (c, r), dtype=x.dtype, device=x.device)
    native.check(lib.ab_transpose2d(x.data_ptr(), y.data_ptr(), batch, r, c, native.dtype_code(x.dtype),
                                    native.stream_ptr()), "ab_transpose2d")
    return y


def embedding(ids: torch.Tensor, tok: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
    """ids [B, L] int64 -> tok[ids] + pos[:L]  [B, L, D]."""
    lib = native.load()
    b, l = ids.shape
    d = tok.shape[1]
    out = torch.empty((b, l, d), dtype=tok.dtype, device=tok.device)
    native.check(lib.ab_embedding(ids.data_ptr(), tok.data_ptr(), pos.data_ptr(), out.data_ptr(), b, l, d,
                                  native.dtype_code(tok.dtype), native.stream_ptr()), "ab_embedding")
    return out
