"""Tensor-parallel GEMMs fused with their collectives over the peer heap (SURVEY §2.8 TP, §8 row F2).

  ag_gemm   column-parallel layer (QKV, FF-in): activations are row-sharded over the subset; each GPU multiplies
            ALL rows by its column slice of W. The all-gather is the kernel's own TMA loads: A tiles are fetched
            from whichever GPU owns the rows (local HBM or a peer over NVLink). No gathered copy exists.
  gemm_rs   row-parallel layer (attn-out, FF-out): each GPU holds a K slice of A and W. K is ALWAYS cut into 8
            canonical shards; a GPU computes its 8/P shards and stores every fp32 partial row straight into the
            workspace of the GPU that owns that row (P2P stores from the tcgen05 epilogue). After a device-side
            barrier each owner adds shards 0..7 in order and applies bias/activation/residual. The sum order never
            depends on P, so outputs are bit-identical for P in {1, 2, 4, 8}.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..ops import native
from ..ops.native import ACT, GemmArgs
from .symm import PeerHeap


class GemmDistArgs(C.Structure):
    _fields_ = [("a_srcs", C.c_void_p * 8), ("nsrc", C.c_int), ("rows_per_src", C.c_int),
                ("ws_peers", C.c_void_p * 8), ("rs_world", C.c_int), ("rs_rank", C.c_int)]


def _lib():
    lib = native.load()
    if not hasattr(lib, "_dist_sig"):
        lib.ab_gemm_dist.restype = C.c_int
        lib.ab_gemm_dist.argtypes = [C.POINTER(GemmArgs), C.POINTER(GemmDistArgs), C.c_int, C.c_void_p]
        lib._dist_sig = True
    return lib


def _base_args(m, n, k, w, out, bias, residual, act, dtype) -> GemmArgs:
    a = GemmArgs()
    a.w, a.out = w.data_ptr(), out.data_ptr()
    a.bias, a.residual = native.ptr(bias), native.ptr(residual)
    a.M, a.N, a.K1, a.K2 = m, n, k, 0
    a.act = ACT[act]
    a.out_scale = 1.0
    a.dtype = native.dtype_code(dtype)
    return a


class AgGemm:
    """All-gather -> GEMM. ``x_slot`` is this rank's row shard in the peer heap (producers write into it)."""

    def __init__(self, heap: PeerHeap, rows_local: int, k: int, dtype=torch.float16):
        assert rows_local % 128 == 0, "row shards must be multiples of 128 rows"
        self.heap, self.rows_local, self.k = heap, rows_local, k
        self.x_slot, self.x_off = heap.alloc((rows_local, k), dtype)

    def __call__(self, w: torch.Tensor, bias: Optional[torch.Tensor] = None, act: Optional[str] = None,
                 block_n: int = 0, sync: bool = True) -> torch.Tensor:
        """w [N_local, K] -> out [world*rows_local, N_local]. x_slot must already hold this rank's rows."""
        heap = self.heap
        if sync:
            heap.barrier()                      # every peer's shard is written
        m = heap.world * self.rows_local
        n = w.shape[0]
        out = torch.empty((m, n), dtype=w.dtype, device=w.device)
        a = _base_args(m, n, self.k, w, out, bias, None, act, w.dtype)
        a.block_n = block_n
        a.splitk = 1
        d = GemmDistArgs()
        if heap.world > 1:
            d.nsrc, d.rows_per_src = heap.world, self.rows_local
            for r in range(heap.world):
                d.a_srcs[r] = heap.peer_ptr(r, self.x_off)
            a.a = None
        else:
            a.a = self.x_slot.data_ptr()
        native.check(_lib().ab_gemm_dist(C.byref(a), C.byref(d), 0, native.stream_ptr()), "ab_gemm_dist(ag)")
        if sync:
            heap.barrier()                      # peers may overwrite their shards only after everyone has read them
        return out


class GemmRs:
    """GEMM -> reduce-scatter with the canonical 8-shard K order."""

    def __init__(self, heap: PeerHeap, m_total: int, n: int):
        assert m_total % heap.world == 0
        self.heap, self.m, self.n = heap, m_total, n
        self.rows_local = m_total // heap.world
        self.ws, self.ws_off = heap.alloc((8, self.rows_local, n), torch.float32)

    def __call__(self, x_kslice: torch.Tensor, w_kslice: torch.Tensor, bias: Optional[torch.Tensor] = None,
                 residual: Optional[torch.Tensor] = None, act: Optional[str] = None, block_n: int = 0) -> torch.Tensor:
        """x [M_total, K_local], w [N, K_local] -> this rank's rows [M_total/world, N]."""
        heap = self.heap
        m, n, k = self.m, self.n, x_kslice.shape[1]
        out = torch.empty((self.rows_local, n), dtype=x_kslice.dtype, device=x_kslice.device)
        a = _base_args(m, n, k, w_kslice, out, bias, residual, act, x_kslice.dtype)
        a.a = x_kslice.data_ptr()
        a.block_n = block_n
        d = GemmDistArgs()
        d.rs_world, d.rs_rank = heap.world, heap.rank
        for r in range(heap.world):
            d.ws_peers[r] = heap.peer_ptr(r, self.ws_off)
        lib = _lib()
        native.check(lib.ab_gemm_dist(C.byref(a), C.byref(d), 1, native.stream_ptr()), "ab_gemm_dist(rs partial)")
        heap.barrier()                          # all partial rows have landed in their owners' workspaces
        native.check(lib.ab_gemm_dist(C.byref(a), C.byref(d), 2, native.stream_ptr()), "ab_gemm_dist(rs reduce)")
        heap.barrier()                          # workspace may be reused
        return out
