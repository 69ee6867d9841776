"""Patch parallelism: ONE image over the GPUs of an NVSwitch box, split into row bands (SURVEY §2.8 "patch
parallelism", §8 row F1; BASELINE config 3). No counterpart in the reference, which runs one GPU per task
(docs/src/pages/mining.mdx:7).

Rank r of a ``PeerHeap`` holds rows [r*H/P, (r+1)*H/P) of every activation of a *sharded* UNet level, channels-last
[B, H/P, W, C]. Three operators cross ranks, each in the canonical order of its single-GPU kernel, so the bytes do not
depend on P:

  GroupNorm       one launch: partial moments of the local canonical shards are stored into every rank's array as
                  16-byte entries that carry the step's tick inside (no fence, no separate flag); every rank polls all 8
                  shards' entries, merges them in shard order and normalises its rows. The rows at the band's edges are
                  also stored into the neighbours' halo slots                     (csrc/norm.cu gn_patch_kernel)
  3x3 conv        reads a [B, H/P + 2, W, C] buffer whose rows 0 and H/P+1 the neighbours wrote; its TMA producer warp
                  waits for their flag before the first load                      (csrc/gemm_tc.cu, args.halo + sync)
  self-attention  K/V projections stay where they were computed; every rank's attention kernel fetches each KV tile
                  from the GPU that owns it by TMA (a tile that straddles two bands in 8-row boxes), waiting for that
                  GPU's flag before its first load from it                         (csrc/attention_v2.cuh SegMaps)

There is no barrier kernel between operators: a hand-off is a flag the producer publishes behind its data (system-scope
release) and the consumer kernel acquires. Flags carry a *tick* that every rank advances once per UNet step, so a
captured CUDA graph replays correctly. Every hand-off has its own buffers and its own flag words ("site"); a buffer is
rewritten one step later, after ~100 all-to-all GroupNorm hand-offs that the reader must have passed, so nothing is
overwritten under a slow reader.

Levels whose height does not split into 8 canonical bands of whole, 8-token-aligned rows (Kandinsky's 12x12 level) are
*replicated*: every rank gathers the full tensor once on the way down, computes the level exactly as one GPU does, and
takes its own band on the way up without communication.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from .. import ops
from ..ops import native
from ..ops.native import ACT, GroupNormArgs
from .symm import PeerHeap

GN_SHARDS = 8
_GN_MAX_SUB = 32


class CopySeg(C.Structure):
    _fields_ = [("dst", C.c_void_p), ("src", C.c_void_p), ("dst_pitch", C.c_longlong), ("src_pitch", C.c_longlong),
                ("width_bytes", C.c_longlong), ("rows", C.c_int)]


class PatchCopyArgs(C.Structure):
    _fields_ = [("seg", CopySeg * 16), ("nseg", C.c_int), ("tick", C.c_void_p), ("wait_site", C.c_void_p),
                ("wait_mask", C.c_uint), ("signal_sites", C.c_void_p * 8), ("world", C.c_int), ("rank", C.c_int),
                ("local_site", C.c_void_p)]


class GnPatchArgs(C.Structure):
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("part_peers", C.c_void_p * 8), ("tick", C.c_void_p),
                ("y_batch_stride", C.c_longlong), ("up_halo", C.c_void_p),
                ("down_halo", C.c_void_p), ("halo_rows", C.c_int), ("done_sites", C.c_void_p * 8),
                ("done_local", C.c_void_p), ("dry", C.c_int)]


class AttnSegsArgs(C.Structure):
    _fields_ = [("q", C.c_void_p), ("out", C.c_void_p), ("batch", C.c_int), ("heads", C.c_int), ("lq", C.c_int),
                ("head_dim", C.c_int), ("q_bs", C.c_longlong), ("q_ls", C.c_longlong), ("o_bs", C.c_longlong),
                ("o_ls", C.c_longlong), ("scale", C.c_float), ("dtype", C.c_int),
                ("prefix", C.c_void_p), ("prefix_len", C.c_int), ("prefix_bs", C.c_longlong), ("prefix_ls", C.c_longlong),
                ("prefix_heads_total", C.c_int), ("prefix_k_head0", C.c_int), ("prefix_v_head0", C.c_int),
                ("pix", C.c_void_p * 8), ("nsrc", C.c_int), ("rows_per_src", C.c_int), ("pix_bs", C.c_longlong),
                ("pix_ls", C.c_longlong), ("heads_total", C.c_int), ("k_head0", C.c_int), ("v_head0", C.c_int),
                ("tick", C.c_void_p), ("wait_site", C.c_void_p), ("self_rank", C.c_int)]


def _lib():
    lib = native.load()
    if not hasattr(lib, "_patch_sig"):
        vp, i, u = C.c_void_p, C.c_int, C.c_uint
        for name, args in (("ab_sync_tick", [vp, vp]),
                           ("ab_sync_signal", [vp, C.POINTER(vp), i, i, vp]),
                           ("ab_sync_wait", [vp, vp, u, vp]),
                           ("ab_patch_copy", [C.POINTER(PatchCopyArgs), vp]),
                           ("ab_groupnorm_patch", [C.POINTER(GroupNormArgs), C.POINTER(GnPatchArgs), vp]),
                           ("ab_flash_attn_segs", [C.POINTER(AttnSegsArgs), vp])):
            fn = getattr(lib, name)
            fn.restype = i
            fn.argtypes = args
        lib._patch_sig = True
    return lib


def flash_attn_segs(q: torch.Tensor, prefix: Optional[torch.Tensor], pix_ptrs: Sequence[int], rows_per_src: int,
                    pix_row_stride: int, pix_batch_stride: int, heads: int, head_dim: int, *, k_head0: int,
                    v_head0: int, heads_total: int, prefix_k_head0: int = 0, prefix_v_head0: int = 0,
                    sync: Optional[Tuple[int, int]] = None, self_rank: int = 0,
                    out: Optional[torch.Tensor] = None, scale: Optional[float] = None) -> torch.Tensor:
    """Attention over the keys [prefix | pixels]. q [B, Lq, H, D] (a strided view is fine); ``prefix`` [B, T, heads_p, D]
    or None; ``pix_ptrs``: device addresses (local or peer-mapped) of every band's fused projection
    [B, rows_per_src, heads_total, D]. K of head h is head ``k_head0 + h`` of those buffers, V is ``v_head0 + h``.
    ``sync`` = (tick_ptr, wait_site_ptr): wait for band s's flag before the first load from it. The tile sequence is
    the same for any number of bands, so the result is bit-identical for P in {1, 2, 4, 8}."""
    lib = _lib()
    b, lq, h, d = q.shape
    assert h == heads and d == head_dim and q.stride(3) == 1 and q.stride(2) == d
    if out is None:
        out = torch.empty((b, lq, h, d), dtype=q.dtype, device=q.device)
    a = AttnSegsArgs()
    a.q, a.out = q.data_ptr(), out.data_ptr()
    a.batch, a.heads, a.lq, a.head_dim = b, h, lq, d
    a.q_bs, a.q_ls, a.o_bs, a.o_ls = q.stride(0), q.stride(1), out.stride(0), out.stride(1)
    a.scale = float(scale) if scale is not None else 0.0
    a.dtype = native.dtype_code(q.dtype)
    if prefix is not None:
        assert prefix.stride(3) == 1 and prefix.stride(2) == d
        a.prefix, a.prefix_len = prefix.data_ptr(), prefix.shape[1]
        a.prefix_bs, a.prefix_ls, a.prefix_heads_total = prefix.stride(0), prefix.stride(1), prefix.shape[2]
        a.prefix_k_head0, a.prefix_v_head0 = prefix_k_head0, prefix_v_head0
    for s, p in enumerate(pix_ptrs):
        a.pix[s] = p
    a.nsrc, a.rows_per_src = len(pix_ptrs), rows_per_src
    a.pix_bs, a.pix_ls, a.heads_total, a.k_head0, a.v_head0 = pix_batch_stride, pix_row_stride, heads_total, k_head0, v_head0
    if sync is not None:
        a.tick, a.wait_site = sync
    a.self_rank = self_rank
    import sys
    _g = sys.modules["arbius_b200.ops.gemm"]
    if _g.RECORD is not None:
        _g.RECORD.append(f"attn:{b}:{h}:{lq}:{(prefix.shape[1] if prefix is not None else 0) + rows_per_src * len(pix_ptrs)}:{d}")
    native.check(lib.ab_flash_attn_segs(C.byref(a), native.stream_ptr()), "ab_flash_attn_segs")
    return out


class Site:
    """16 u32 at one heap offset on every rank: word r = rank r's flag, word 8 = local arrival counter."""

    def __init__(self, heap: PeerHeap):
        self.t, self.off = heap.alloc((16,), torch.int32)
        self.local = heap.peer_ptr(heap.rank, self.off)
        self.peers = [heap.peer_ptr(r, self.off) for r in range(heap.world)]


def shardable(h: int, w: int) -> bool:
    """A level splits into canonical row bands when its 8 shards are whole rows of 8-token-aligned length (the
    attention kernel assembles a straddling KV tile from 8-row boxes). Independent of the GPU count."""
    return h % GN_SHARDS == 0 and ((h // GN_SHARDS) * w) % 8 == 0


class PatchContext:
    """Per-(heap, batch, dtype) runtime: tick, flag sites, and the cross-rank operators."""

    def __init__(self, heap: PeerHeap, dtype=torch.float16):
        assert heap.world in (2, 4, 8), "patch parallelism runs on 2, 4 or 8 GPUs"
        self.heap, self.world, self.rank, self.dtype = heap, heap.world, heap.rank, dtype
        self.esz = torch.empty((), dtype=dtype).element_size()
        self.tick_t = torch.zeros(4, dtype=torch.int32, device="cuda")
        self.tick_ptr = self.tick_t.data_ptr()
        self.up = self.rank - 1 if self.rank > 0 else None
        self.down = self.rank + 1 if self.rank < self.world - 1 else None
        self.all_mask = ((1 << self.world) - 1) & ~(1 << self.rank)
        self.nb_mask = (0 if self.up is None else 1 << self.up) | (0 if self.down is None else 1 << self.down)
        self.launches = 0

    @property
    def dry(self) -> bool:
        return bool(self.heap.dry)

    # ---- primitives -------------------------------------------------------------------------------------------
    def tick(self) -> None:
        """Once per UNet step, on every rank, before the step's first hand-off."""
        native.check(_lib().ab_sync_tick(self.tick_ptr, native.stream_ptr()), "ab_sync_tick")

    def site(self) -> Site:
        return Site(self.heap)

    def signal(self, site: Site, ranks: Optional[Sequence[int]] = None) -> None:
        if self.dry:
            return
        arr = (C.c_void_p * 8)()
        for r in (ranks if ranks is not None else range(self.world)):
            if r != self.rank:
                arr[r] = site.peers[r]
        native.check(_lib().ab_sync_signal(self.tick_ptr, arr, self.world, self.rank, native.stream_ptr()), "ab_sync_signal")

    def wait(self, site: Site, mask: int) -> None:
        if self.dry or not mask:
            return
        native.check(_lib().ab_sync_wait(self.tick_ptr, site.local, mask, native.stream_ptr()), "ab_sync_wait")

    def conv_sync(self, site: Site):
        """``sync=`` argument of ops.conv3x3: wait for the neighbours' halo rows inside the conv kernel."""
        if self.dry or not self.nb_mask:
            return None
        return (self.tick_ptr, site.local, self.nb_mask, 0 if self.up is None else 1 << self.up)

    def copy(self, segs: List[Tuple[int, int, int, int, int, int]], wait: Optional[Tuple[Site, int]] = None,
             signal: Optional[Tuple[Site, Sequence[int]]] = None) -> None:
        """segs: (dst_ptr, src_ptr, dst_pitch, src_pitch, width_bytes, rows). wait=(site, mask): every CTA waits first;
        signal=(site, ranks): the last CTA publishes the tick to those ranks."""
        a = PatchCopyArgs()
        assert len(segs) <= 16
        for i, (d, s, dp, sp, wb, rows) in enumerate(segs):
            a.seg[i].dst, a.seg[i].src, a.seg[i].dst_pitch, a.seg[i].src_pitch = d, s, dp, sp
            a.seg[i].width_bytes, a.seg[i].rows = wb, rows
        a.nseg, a.tick, a.world, a.rank = len(segs), self.tick_ptr, self.world, self.rank
        if wait is not None and not self.dry and wait[1]:
            a.wait_site, a.wait_mask = wait[0].local, wait[1]
        if signal is not None and not self.dry:
            for r in signal[1]:
                if r != self.rank:
                    a.signal_sites[r] = signal[0].peers[r]
            a.local_site = signal[0].local
        native.check(_lib().ab_patch_copy(C.byref(a), native.stream_ptr()), "ab_patch_copy")

    # ---- buffers ------------------------------------------------------------------------------------------------
    def halo_buffer(self, b: int, hl: int, w: int, c: int):
        """[B, hl + 2, W, C] in the heap (zero-filled: the image's top and bottom halo rows stay zero for good)."""
        return self.heap.alloc((b, hl + 2, w, c), self.dtype)

    def neighbours(self) -> List[int]:
        return [r for r in (self.up, self.down) if r is not None]


class PatchGroupNorm:
    """One GroupNorm call site of a sharded level: its partial-moment array, its stats site and (when a 3x3 conv
    follows) its halo buffer and halo site. ``__call__`` is ONE kernel launch."""

    def __init__(self, ctx: PatchContext, b: int, hl: int, w: int, c: int, groups: int, halo: bool,
                 done_to_all: bool = False):
        self.ctx, self.b, self.hl, self.w, self.c, self.groups = ctx, b, hl, w, c, groups
        heap = ctx.heap
        nbytes = b * GN_SHARDS * _GN_MAX_SUB * max(groups, 32) * 16
        self.part, self.part_off = heap.alloc((nbytes,), torch.uint8)
        self.halo = halo
        self.done = ctx.site() if (halo or done_to_all) else None
        self.done_to_all = done_to_all
        if halo:
            self.buf, self.buf_off = ctx.halo_buffer(b, hl, w, c)
        else:
            self.buf, self.buf_off = heap.alloc((b, hl, w, c), ctx.dtype) if done_to_all else (None, 0)

    def __call__(self, x: torch.Tensor, gamma, beta, eps: float, act: Optional[str], x2: Optional[torch.Tensor] = None,
                 scale: Optional[torch.Tensor] = None, shift: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x [B, hl, W, C1] (+ x2) local band -> normalised band. With ``halo`` the result is the [B, hl+2, W, C] halo
        buffer (interior written here, edge rows pushed to / by the neighbours; pass ``ctx.conv_sync(self.done)`` to
        the conv that reads it)."""
        ctx, heap = self.ctx, self.ctx.heap
        n, c1 = x.shape[0], x.shape[-1]
        c2 = x2.shape[-1] if x2 is not None else 0
        rows = self.hl * self.w
        assert x.is_contiguous() and (x2 is None or x2.is_contiguous()) and c1 + c2 == self.c and n == self.b
        assert x.numel() == n * rows * c1
        a = GroupNormArgs()
        a.x, a.x2 = x.data_ptr(), native.ptr(x2)
        a.gamma, a.beta = native.ptr(gamma), native.ptr(beta)
        a.n, a.hw, a.c1, a.c2, a.groups = n, rows, c1, c2, self.groups
        a.eps, a.act, a.dtype = eps, ACT[act], native.dtype_code(x.dtype)
        a.scale, a.shift = native.ptr(scale), native.ptr(shift)
        if scale is not None and scale.dim() == 2:
            a.ss_stride = scale.stride(0)
        pa = GnPatchArgs()
        pa.world, pa.rank, pa.tick, pa.dry = ctx.world, ctx.rank, ctx.tick_ptr, int(ctx.dry)
        for r in range(ctx.world):
            pa.part_peers[r] = heap.peer_ptr(r, self.part_off)
        row_b = self.w * self.c * ctx.esz
        if self.halo:
            y = self.buf
            a.y = y[:, 1].data_ptr()
            pa.y_batch_stride = (self.hl + 2) * self.w * self.c
            pa.halo_rows = self.w
            if ctx.up is not None:       # my first image row -> the upper neighbour's bottom halo slot (row hl+1)
                pa.up_halo = heap.peer_ptr(ctx.up, self.buf_off) + (self.hl + 1) * row_b
                pa.done_sites[ctx.up] = self.done.peers[ctx.up]
            if ctx.down is not None:     # my last image row -> the lower neighbour's top halo slot (row 0)
                pa.down_halo = heap.peer_ptr(ctx.down, self.buf_off)
                pa.done_sites[ctx.down] = self.done.peers[ctx.down]
            pa.done_local = self.done.local
        else:
            y = self.buf if self.buf is not None else torch.empty((n, self.hl, self.w, self.c), dtype=x.dtype, device=x.device)
            a.y = y.data_ptr()
            if self.done_to_all:
                for r in range(ctx.world):
                    if r != ctx.rank:
                        pa.done_sites[r] = self.done.peers[r]
                pa.done_local = self.done.local
        native.check(_lib().ab_groupnorm_patch(C.byref(a), C.byref(pa), native.stream_ptr()), "ab_groupnorm_patch")
        return y


class HaloExchange:
    """A [B, hl+2, W, C] heap buffer whose interior an elementwise op fills (a resampled band); ``push`` sends the edge
    rows to the neighbours and publishes the flag the following conv waits on."""

    def __init__(self, ctx: PatchContext, b: int, hl: int, w: int, c: int):
        self.ctx, self.b, self.hl, self.w, self.c = ctx, b, hl, w, c
        self.buf, self.off = ctx.halo_buffer(b, hl, w, c)
        self.site = ctx.site()

    def interior(self, i: int) -> torch.Tensor:
        return self.buf[i, 1:self.hl + 1]

    def push(self) -> None:
        ctx, heap = self.ctx, self.ctx.heap
        row_b = self.w * self.c * ctx.esz
        pitch = (self.hl + 2) * row_b
        base = self.buf.data_ptr()
        segs = []
        if ctx.up is not None:
            segs.append((heap.peer_ptr(ctx.up, self.off) + (self.hl + 1) * row_b, base + row_b, pitch, pitch, row_b, self.b))
        if ctx.down is not None:
            segs.append((heap.peer_ptr(ctx.down, self.off), base + self.hl * row_b, pitch, pitch, row_b, self.b))
        ctx.copy(segs, signal=(self.site, ctx.neighbours()))


class BandGather:
    """All-gather of row bands into a full [B, H, W, C] tensor on every rank (level too small to split; the UNet's
    output). ``push(band)`` stores the local band into every rank's full buffer and publishes; ``full()`` waits."""

    def __init__(self, ctx: PatchContext, b: int, h: int, w: int, c: int, dtype=None):
        self.ctx, self.b, self.h, self.w, self.c = ctx, b, h, w, c
        self.dtype = dtype or ctx.dtype
        self.esz = torch.empty((), dtype=self.dtype).element_size()
        self.buf, self.off = ctx.heap.alloc((b, h, w, c), self.dtype)
        self.site = ctx.site()
        self.hl = h // ctx.world

    def push(self, band: torch.Tensor, src_batch_pitch: Optional[int] = None, src_ptr: Optional[int] = None) -> None:
        """band [B, hl, W, C] contiguous (or ``src_ptr`` / ``src_batch_pitch`` in bytes for the interior of a halo buffer)."""
        ctx, heap = self.ctx, self.ctx.heap
        band_b = self.hl * self.w * self.c * self.esz
        dpitch = self.h * self.w * self.c * self.esz
        sp = src_batch_pitch if src_batch_pitch is not None else band_b
        src = src_ptr if src_ptr is not None else band.data_ptr()
        segs = [(heap.peer_ptr(r, self.off) + ctx.rank * band_b, src, dpitch, sp, band_b, self.b) for r in range(ctx.world)]
        ctx.copy(segs, signal=(self.site, list(range(ctx.world))))

    def full(self) -> torch.Tensor:
        self.ctx.wait(self.site, self.ctx.all_mask)
        return self.buf
