"""Patch-parallel self-attention: the displaced-activation all-gather fused into the K/V loads.

Each GPU of a subset owns a contiguous band of tokens (row band of the latent image). Its K and V projections are
written into the peer heap; queries stay local. ``patch_flash_attn`` then launches ONE attention kernel per GPU
whose K/V tiles are fetched by TMA straight from whichever GPU owns them (local HBM or a peer over NVLink), in the
global tile order — there is no all-gather kernel, no staging copy, and the result is bit-identical to the
single-GPU kernel for any subset size (SURVEY §2.8 patch parallelism, §8 row F1).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..ops import native
from ..ops.native import AttnArgs
from .symm import PeerHeap


def _sig(lib):
    if not hasattr(lib, "_attn_multi_sig"):
        lib.ab_flash_attn_multi.restype = C.c_int
        lib.ab_flash_attn_multi.argtypes = [C.POINTER(AttnArgs), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int,
                                            C.c_int, C.c_void_p]
        lib._attn_multi_sig = True
    return lib


def flash_attn_multi(q: torch.Tensor, k_ptrs, v_ptrs, lk_per_src: int, kv_row_stride: int, kv_batch_stride: int,
                     scale: Optional[float] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q [B, Lq, H, D] local; k_ptrs/v_ptrs: device addresses of each source's [B, lk_per_src, H, D] buffer."""
    lib = _sig(native.load())
    b, lq, h, d = q.shape
    if out is None:
        out = torch.empty((b, lq, h, d), dtype=q.dtype, device=q.device)
    a = AttnArgs()
    a.q, a.out = q.data_ptr(), out.data_ptr()
    a.k = a.v = None
    a.batch, a.heads, a.lq, a.lk, a.head_dim = b, h, lq, lk_per_src * len(k_ptrs), d
    a.q_bs, a.q_ls = q.stride(0), q.stride(1)
    a.k_bs = a.v_bs = kv_batch_stride
    a.k_ls = a.v_ls = kv_row_stride
    a.o_bs, a.o_ls = out.stride(0), out.stride(1)
    a.scale = float(scale) if scale is not None else 0.0
    a.causal = 0
    a.dtype = native.dtype_code(q.dtype)
    n = len(k_ptrs)
    ka = (C.c_void_p * n)(*k_ptrs)
    va = (C.c_void_p * n)(*v_ptrs)
    native.check(lib.ab_flash_attn_multi(C.byref(a), ka, va, n, lk_per_src, native.stream_ptr()), "ab_flash_attn_multi")
    return out


class PatchAttention:
    """Owns the K/V slots in the peer heap for one attention layer shape."""

    def __init__(self, heap: PeerHeap, batch: int, tokens_local: int, heads: int, head_dim: int, dtype=torch.float16):
        self.heap = heap
        self.shape = (batch, tokens_local, heads, head_dim)
        self.k, self.k_off = heap.alloc(self.shape, dtype)
        self.v, self.v_off = heap.alloc(self.shape, dtype)

    def __call__(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, scale: Optional[float] = None) -> torch.Tensor:
        """q, k, v: this rank's [B, L_local, H, D] projections. Returns this rank's attention output."""
        self.k.copy_(k)         # in the fused model the QKV GEMM writes here directly (output pointer = heap slot)
        self.v.copy_(v)
        self.heap.barrier()     # peers' K/V are complete and visible
        b, l, h, d = self.shape
        kp = [self.heap.peer_ptr(r, self.k_off) for r in range(self.heap.world)]
        vp = [self.heap.peer_ptr(r, self.v_off) for r in range(self.heap.world)]
        out = flash_attn_multi(q, kp, vp, l, h * d, l * h * d, scale)
        self.heap.barrier()     # nobody overwrites K/V while a peer is still reading them
        return out
