"""Flash-attention wrapper (csrc/attention.cu). Tensors are token-major [B, L, H, D] *views*:
the last dim must be contiguous, head stride must equal D, batch/row strides are free — so q, k, v
can be column slices of one fused QKV projection output and no permute copy is ever made."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import native
from .native import AttnArgs


def flash_attn(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, *, scale: Optional[float] = None,
               causal: bool = False, out: Optional[torch.Tensor] = None, long_queries: bool = False) -> torch.Tensor:
    """``long_queries``: choose the kernel as for more than 128 query rows per batch element even if this call has
    fewer — for a rank that holds a slice of a longer query sequence, so the bytes do not depend on the split."""
    lib = native.load()
    b, lq, h, d = q.shape
    lk = k.shape[1]
    for t in (q, k, v):
        assert t.stride(3) == 1 and t.stride(2) == d, "need [B, L, H, D] with contiguous (H, D)"
    if out is None:
        out = torch.empty((b, lq, h, d), dtype=q.dtype, device=q.device)
    a = AttnArgs()
    a.q, a.k, a.v, a.out = q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr()
    a.batch, a.heads, a.lq, a.lk, a.head_dim = b, h, lq, lk, d
    a.q_bs, a.q_ls = q.stride(0), q.stride(1)
    a.k_bs, a.k_ls = k.stride(0), k.stride(1)
    a.v_bs, a.v_ls = v.stride(0), v.stride(1)
    a.o_bs, a.o_ls = out.stride(0), out.stride(1)
    a.scale = float(scale) if scale is not None else 0.0
    a.causal = int(causal) | (2 if long_queries else 0)
    a.dtype = native.dtype_code(q.dtype)
    import sys
    _g = sys.modules["arbius_b200.ops.gemm"]
    if _g.RECORD is not None:
        _g.RECORD.append(f"attn:{b}:{h}:{lq}:{lk}:{d}")
    native.check(lib.ab_flash_attn(C.byref(a), native.stream_ptr()), "ab_flash_attn")
    return out
