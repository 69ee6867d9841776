"""Sequence (frame) parallelism for the text-to-video UNet3D over the GPUs of one NVSwitch box (SURVEY §2.8 row
"Sequence parallelism / Ulysses"; no counterpart in the reference, which runs one GPU per task).

Rank r of a ``PeerHeap`` holds frames [r*F/P, (r+1)*F/P) of every activation, channels-last [B, F/P, H, W, C]. Spatial
layers (resnets, spatial transformers, up/down-sampling, VAE) are per-frame and run unchanged on the local frames; the
three temporal operators cross ranks through peer memory, each in the canonical order of the single-GPU kernel so the
bytes do not depend on P:

  GroupNorm over (F, H, W)   each rank reduces its own canonical row shards (8 per sample, 8/P per rank), writes the
                             partial moments to its heap, barrier, every rank merges all partials (peer loads) in
                             shard order and normalises its rows            (csrc/norm.cu ab_groupnorm_sharded_*)
  temporal Conv3d (3,1,1)    the normalised frames go to a padded [B, F/P + 2, HW, C] heap buffer, barrier, the two
                             neighbour frames are copied in from the ranks next door (zeros at the clip's ends) and
                             the implicit-GEMM conv reads the halo rows as data  (csrc/gemm_tc.cu, args.halo)
  temporal attention         fused all-to-all: the QKV projection lands in the heap, barrier, every rank takes a pixel
                             band, pulls that band's Q/K/V frames from ALL ranks by TMA, and stores each frame's
                             output rows to the rank that owns the frame     (csrc/temporal_attn.cu, pix_begin/count)

Remote tiles are fetched per CTA with plain (non-multicast) TMA loads. Every barrier is the graph-replayable device
barrier of ``PeerHeap``; buffers that peers read alternate by parity where the next write could overtake a slow reader.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import ops
from ..ops import native
from ..ops.native import ACT, GroupNormArgs
from ..ops.temporal import temporal_attn_ptrs
from .symm import PeerHeap

GN_SHARDS = 8
_GN_MAX_SUB = 32


def _lib():
    lib = native.load()
    if not hasattr(lib, "_fs_sig"):
        vp = C.c_void_p
        lib.ab_groupnorm_sharded_stats.restype = C.c_int
        lib.ab_groupnorm_sharded_stats.argtypes = [C.POINTER(GroupNormArgs), C.c_int, C.c_int, vp, vp]
        lib.ab_groupnorm_sharded_apply.restype = C.c_int
        lib.ab_groupnorm_sharded_apply.argtypes = [C.POINTER(GroupNormArgs), C.c_int, C.POINTER(vp), C.c_longlong, vp]
        lib.ab_peer_copy2d.restype = C.c_int
        lib.ab_peer_copy2d.argtypes = [vp, C.c_longlong, vp, C.c_longlong, C.c_longlong, C.c_int, vp]
        lib._fs_sig = True
    return lib


class FrameShard:
    """Heap layout + the three cross-rank temporal operators for one (batch, frames, largest activation) budget."""

    def __init__(self, heap: PeerHeap, batch: int, frames_local: int, max_hw: int, max_c: int, dtype=torch.float16,
                 groups: int = 32):
        self.heap, self.world, self.rank = heap, heap.world, heap.rank
        assert self.world in (2, 4, 8), "frame sharding needs 2, 4 or 8 ranks"
        self.dtype, self.batch, self.fl = dtype, batch, frames_local
        esz = torch.empty((), dtype=dtype).element_size()
        self.esz = esz
        # partial GroupNorm moments, two parities: [n][8][sub][groups] float4
        part_bytes = batch * GN_SHARDS * _GN_MAX_SUB * max(groups, 32) * 16
        self.part = [heap.alloc((part_bytes,), torch.uint8) for _ in range(2)]
        # padded conv inputs, two parities; QKV and attention output, one each
        act = batch * max_hw * max_c
        self.pad = [heap.alloc((act * (frames_local + 2),), dtype) for _ in range(2)]
        self.qkv = heap.alloc((act * frames_local * 3,), dtype)
        self.att = heap.alloc((act * frames_local,), dtype)
        self._gn_par = 0
        self.barriers = 0

    @staticmethod
    def heap_bytes(batch: int, frames_local: int, max_hw: int, max_c: int, esz: int = 2, groups: int = 32) -> int:
        act = batch * max_hw * max_c * esz
        return (2 * batch * GN_SHARDS * _GN_MAX_SUB * max(groups, 32) * 16 + 2 * act * (frames_local + 2)
                + 4 * act * frames_local + (1 << 20))

    def barrier(self) -> None:
        self.heap.barrier_replayable()
        self.barriers += 1

    # ---- GroupNorm over all frames -------------------------------------------------------------------------------
    def groupnorm(self, x: torch.Tensor, gamma, beta, groups: int, eps: float, act: Optional[str],
                  padded_parity: Optional[int] = None) -> torch.Tensor:
        """x [B, F_local*HW, C] (local rows of every sample) -> same shape, statistics over ALL ranks' rows. With
        ``padded_parity`` the result is written to the interior of that parity's padded conv buffer and the padded
        [B, F_local + 2, HW, C] view is returned (``x`` must then be [B, F_local, HW, C])."""
        lib = _lib()
        n, c = x.shape[0], x.shape[-1]
        rows = x.numel() // (n * c)
        assert x.is_contiguous() and rows % (GN_SHARDS // self.world) == 0
        par = self._gn_par
        self._gn_par ^= 1
        part, part_off = self.part[par]
        a = GroupNormArgs()
        a.x, a.gamma, a.beta = x.data_ptr(), native.ptr(gamma), native.ptr(beta)
        a.partials = ops.norm._gn_workspace(n, groups, x.device).data_ptr()
        a.n, a.hw, a.c1, a.c2, a.groups = n, rows, c, 0, groups
        a.eps, a.act, a.dtype = eps, ACT[act], native.dtype_code(x.dtype)
        native.check(lib.ab_groupnorm_sharded_stats(C.byref(a), self.world, self.rank, part.data_ptr(),
                                                    native.stream_ptr()), "ab_groupnorm_sharded_stats")
        self.barrier()                                    # every rank's partial moments are written
        y_bs = 0
        if padded_parity is None:
            y = torch.empty_like(x)
            a.y = y.data_ptr()
        else:
            fl = self.fl
            hw = rows // fl
            buf, _ = self.pad[padded_parity]
            y = buf[:n * (fl + 2) * hw * c].view(n, fl + 2, hw, c)
            a.y = y[:, 1].data_ptr()                      # first interior frame of sample 0
            y_bs = (fl + 2) * hw * c
        srcs = (C.c_void_p * 8)(*[self.heap.peer_ptr(r, part_off) for r in range(self.world)] + [None] * (8 - self.world))
        native.check(lib.ab_groupnorm_sharded_apply(C.byref(a), self.world, srcs, y_bs, native.stream_ptr()),
                     "ab_groupnorm_sharded_apply")
        return y

    # ---- temporal conv with neighbour frames ---------------------------------------------------------------------
    def exchange_halos(self, padded: torch.Tensor, parity: int) -> None:
        """Fill frames 0 and F_local+1 of ``padded`` [B, F_local+2, HW, C] from the neighbours' interiors."""
        lib = _lib()
        n, fp, hw, c = padded.shape
        fl = fp - 2
        self.barrier()                                    # every rank's interior frames are written
        frame_b = hw * c * self.esz
        pitch = fp * frame_b
        _, off = self.pad[parity]
        st = native.stream_ptr()
        prev_src = self.heap.peer_ptr(self.rank - 1, off) + fl * frame_b if self.rank > 0 else None
        next_src = self.heap.peer_ptr(self.rank + 1, off) + 1 * frame_b if self.rank < self.world - 1 else None
        native.check(lib.ab_peer_copy2d(padded.data_ptr(), pitch, prev_src, pitch, frame_b, n, st), "ab_peer_copy2d")
        native.check(lib.ab_peer_copy2d(padded.data_ptr() + (fl + 1) * frame_b, pitch, next_src, pitch, frame_b, n, st),
                     "ab_peer_copy2d")

    # ---- temporal attention as a fused all-to-all ----------------------------------------------------------------
    def qkv_buffer(self, b: int, hw: int, c3: int) -> torch.Tensor:
        buf, _ = self.qkv
        return buf[:b * self.fl * hw * c3].view(b * self.fl * hw, c3)

    def temporal_attention(self, b: int, hw: int, heads: int, c: int) -> torch.Tensor:
        """The QKV projection of the local frames is in ``qkv_buffer`` ([B*F_local*HW, 3C]); returns the attention
        output of the local frames [B*F_local*HW, C] (rows written by whichever rank owned the pixel band)."""
        fl, w, r = self.fl, self.world, self.rank
        _, qoff = self.qkv
        abuf, aoff = self.att
        self.barrier()                                    # every rank's Q/K/V are written
        q = [self.heap.peer_ptr(s, qoff) for s in range(w)]
        k = [p + c * self.esz for p in q]
        v = [p + 2 * c * self.esz for p in q]
        o = [self.heap.peer_ptr(s, aoff) for s in range(w)]
        # bands are whole tiles of the single-GPU launch (128 // F pixels each): a pixel keeps its slot in its tile, so
        # the tensor core sums its keys in the same positions and the output bits do not depend on the rank count
        px = 128 // (fl * w)
        tiles = (hw + px - 1) // px
        per_rank = (tiles + w - 1) // w
        p0, p1 = min(hw, r * per_rank * px), min(hw, (r + 1) * per_rank * px)
        if p1 > p0:
            temporal_attn_ptrs(q, k, v, o, b, fl * w, hw, heads, c // heads, 3 * c, c, self.dtype,
                               pix_begin=p0, pix_count=p1 - p0)
        self.barrier()                                    # every band of the local frames' output has arrived
        return abuf[:b * fl * hw * c].view(b * fl * hw, c)
