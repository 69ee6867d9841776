"""Temporal attention wrapper (csrc/temporal_attn.cu): pixels are the batch, frames the sequence."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import native


class TemporalAttnArgs(C.Structure):
    _fields_ = [("q", C.c_void_p * 8), ("k", C.c_void_p * 8), ("v", C.c_void_p * 8), ("out", C.c_void_p * 8),
                ("nsrc", C.c_int), ("batch", C.c_int), ("heads", C.c_int), ("frames", C.c_int), ("hw", C.c_int),
                ("head_dim", C.c_int), ("row_stride", C.c_longlong), ("out_row_stride", C.c_longlong),
                ("scale", C.c_float), ("dtype", C.c_int), ("pix_begin", C.c_int), ("pix_count", C.c_int)]


def _lib():
    lib = native.load()
    if not hasattr(lib, "_ta_sig"):
        lib.ab_temporal_attn.restype = C.c_int
        lib.ab_temporal_attn.argtypes = [C.POINTER(TemporalAttnArgs), C.c_void_p]
        lib._ta_sig = True
    return lib


def temporal_attn_ptrs(q_ptrs: Sequence[int], k_ptrs: Sequence[int], v_ptrs: Sequence[int], out_ptrs: Sequence[int],
                       batch: int, frames: int, hw: int, heads: int, head_dim: int, row_stride: int,
                       out_row_stride: int, dtype: torch.dtype, scale: Optional[float] = None,
                       pix_begin: int = 0, pix_count: int = 0) -> None:
    """Multi-source form: source s holds frames [s*F/n, (s+1)*F/n) as [B, F/n, HW, row_stride]; ``frames`` is the total.
    ``pix_count`` > 0 restricts the launch to a pixel band (the frame-sharded UNet3D: one band per GPU)."""
    a = TemporalAttnArgs()
    a.pix_begin, a.pix_count = pix_begin, pix_count
    n = len(q_ptrs)
    for s in range(n):
        a.q[s], a.k[s], a.v[s], a.out[s] = q_ptrs[s], k_ptrs[s], v_ptrs[s], out_ptrs[s]
    a.nsrc, a.batch, a.heads, a.frames, a.hw, a.head_dim = n, batch, heads, frames, hw, head_dim
    a.row_stride, a.out_row_stride = row_stride, out_row_stride
    a.scale = float(scale) if scale is not None else 0.0
    a.dtype = native.dtype_code(dtype)
    native.check(_lib().ab_temporal_attn(C.byref(a), native.stream_ptr()), "ab_temporal_attn")


def temporal_attn(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, scale: Optional[float] = None,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q, k, v: [B, F, HW, heads*D] views with a common token stride (e.g. slices of a fused QKV output)."""
    b, f, hw, c = q.shape
    d = c // heads
    assert q.stride(3) == 1 and q.stride(2) == k.stride(2) == v.stride(2)
    assert q.stride(1) == hw * q.stride(2) and q.stride(0) == f * hw * q.stride(2), "token-major layout required"
    if out is None:
        out = torch.empty((b, f, hw, c), dtype=q.dtype, device=q.device)
    temporal_attn_ptrs([q.data_ptr()], [k.data_ptr()], [v.data_ptr()], [out.data_ptr()], b, f, hw, heads, d,
                       q.stride(2), out.stride(2), q.dtype, scale)
    return out
