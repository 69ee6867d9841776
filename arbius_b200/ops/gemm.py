"""GEMM / conv wrappers over ``ab_gemm`` and ``ab_conv_direct`` (csrc/gemm_tc.cu, conv_direct.cu).

Layouts (all channels-last):
  linear   x [M, K]            w [N, K]                 -> [M, N]
  conv3x3  x [B, H, W, C]      w [Cout, 3, 3, C] packed  -> [B, H, W, Cout]
Weights arrive in their upstream layout (``nn.Linear`` [N, K], ``nn.Conv2d``
[Cout, Cin, kh, kw]) and are repacked once at load time by ``pack_conv_weight`` /
``pack_geglu_weight``; the checkpoint key layout on disk is untouched.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import json
import threading
from pathlib import Path
from typing import Optional

import torch

from . import native
from .native import ACT, GemmArgs

_workspaces: dict = {}

# Shape -> (block_n, splitk) choices measured on a B200 by scripts/tune_gemm.py. The table is part of the
# source tree, so every box picks the same tiling for the same shape (bit-identical results everywhere).
_TUNED_PATH = Path(__file__).with_name("tuned_sm100.json")
_tuned: Optional[dict] = None
RECORD: Optional[list] = None      # set to a list to log every (key, call kwargs) that goes through here
# Canonical-batch multiplier: when one task's batch is split over GPUs (classifier-free-guidance parallelism: each GPU
# runs one of the two UNet branches) every tiling decision that changes the summation order (split-K) must be the one
# the undivided batch would get, or the bytes would depend on the GPU count. The models set this around their calls.
# A Fraction 1/n is set while n independent tasks share one batch: each task then gets the choices it gets alone.
# The multiplier is per host thread: the miner drives one thread per GPU in one process, and a thread that warms up a
# pair / frame / batched predictor must not change the split-K another thread bakes into the graph it is capturing.
_scale_tls = threading.local()


def get_batch_scale():
    return getattr(_scale_tls, "value", 1)


@contextlib.contextmanager
def batch_scale(value, retile: bool = False):
    """``with batch_scale(v):`` — the canonical-batch multiplier for the GEMM/conv calls of THIS thread.
    ``retile``: the calls hold 1/v of the canonical rows for good (row bands of patch parallelism), so the tile WIDTH —
    which never changes a sum's order — is chosen for the rows really present instead of the canonical shape's tuned
    width (a band needs narrower tiles to fill the SMs); split-K stays the canonical shape's."""
    prev = get_batch_scale(), getattr(_scale_tls, "retile", False)
    _scale_tls.value, _scale_tls.retile = value, retile
    try:
        yield
    finally:
        _scale_tls.value, _scale_tls.retile = prev


def _canon(m: int) -> int:
    return int(m * get_batch_scale())


def _tuned_pair(key: str, actual_key: str):
    """(block_n, splitk) for a call: split-K — the only choice that changes the summation order — always comes from the
    canonical shape's entry; the tile width is free, so a batch of several tasks takes the entry of its real shape if
    there is one and otherwise lets the library size the tile for the rows it really has (block_n 0)."""
    cfg = tuned_config(key)
    if cfg is not None and getattr(_scale_tls, "retile", False) and key != actual_key:
        own = tuned_config(actual_key)
        return (own[0] if own is not None else 0, cfg[1])
    if cfg is None or get_batch_scale() >= 1 or key == actual_key or _BATCH_BN == "canon":
        return cfg
    own = tuned_config(actual_key)
    if own is None and _BATCH_BN == "small" and cfg[0] > 96:
        return cfg                                      # wide tiles were not chosen for want of rows: keep them
    return (own[0] if own is not None else 0, cfg[1])


import os as _os

_BATCH_BN = _os.environ.get("AB_BATCH_BN", "auto")


def tuned_config(key: str):
    global _tuned
    if _tuned is None:
        _tuned = json.loads(_TUNED_PATH.read_text()) if _TUNED_PATH.exists() else {}
    return _tuned.get(key)


_retired_workspaces: list = []
_ws_lock = threading.Lock()


def _workspace(nbytes: int, device) -> torch.Tensor:
    """Per-device split-K scratch, grown on demand (never shrinks, reused stream-ordered). A buffer that is outgrown is
    kept alive for the life of the process: CUDA graphs captured earlier hold its raw address, and freeing it would let
    their replays write fp32 partials into memory the allocator has handed to another tensor."""
    key = (device.index if device.index is not None else torch.cuda.current_device())
    with _ws_lock:
        ws = _workspaces.get(key)
        if ws is None or ws.numel() < nbytes:
            if ws is not None:
                _retired_workspaces.append(ws)
            ws = torch.empty(max(nbytes, 1 << 26), dtype=torch.uint8, device=device)
            _workspaces[key] = ws
        return ws


def pack_conv_weight(w: torch.Tensor) -> torch.Tensor:
    """[Cout, Cin, kh, kw] (torch) -> [Cout, kh*kw*Cin] with K ordered (kh, kw, cin)."""
    co, ci, kh, kw = w.shape
    return w.permute(0, 2, 3, 1).reshape(co, kh * kw * ci).contiguous()


def pack_geglu_weight(w: torch.Tensor, b: Optional[torch.Tensor]):
    """Interleave value/gate rows in blocks of 16 so one 32-column accumulator chunk holds
    16 values and their 16 gates. w is the upstream ``GEGLU.proj`` weight [2*inner, K]
    (first half values, second half gates)."""
    two_inner, k = w.shape
    inner = two_inner // 2
    assert inner % 16 == 0, "GEGLU inner dim must be a multiple of 16"
    val = w[:inner].reshape(inner // 16, 16, k)
    gate = w[inner:].reshape(inner // 16, 16, k)
    wp = torch.stack([val, gate], dim=1).reshape(two_inner, k).contiguous()
    bp = None
    if b is not None:
        bp = torch.stack([b[:inner].reshape(-1, 16), b[inner:].reshape(-1, 16)], dim=1).reshape(two_inner).contiguous()
    return wp, bp


def auto_splitk(m: int, n: int, k_chunks: int, m_tiles: Optional[int] = None) -> int:
    """Fixed, shape-only split-K choice (never depends on GPU count or load)."""
    m_tiles = m_tiles if m_tiles is not None else (m + 127) // 128
    tiles = m_tiles * ((n + 127) // 128)
    if tiles >= 96 or k_chunks < 16:
        return 1
    s = min(8, max(1, 148 // tiles), k_chunks // 8)
    return max(1, s)


def gemm(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
         x2: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
         rowbias: Optional[torch.Tensor] = None, rows_per_group: int = 0,
         act: Optional[str] = None, geglu: bool = False, out_scale: float = 1.0,
         out: Optional[torch.Tensor] = None, splitk: int = 0, block_n: int = 0) -> torch.Tensor:
    """out = epilogue(cat[x, x2] @ w.T). x [..., K1], w [N, K1+K2]."""
    lib = native.load()
    lead = x.shape[:-1]
    k1 = x.shape[-1]
    xm = x.reshape(-1, k1)
    assert xm.is_contiguous() and w.is_contiguous()
    m = xm.shape[0]
    n = w.shape[0]
    k2 = 0
    x2m = None
    if x2 is not None:
        k2 = x2.shape[-1]
        x2m = x2.reshape(-1, k2)
        assert x2m.is_contiguous() and x2m.shape[0] == m
    assert w.shape[1] == k1 + k2, (w.shape, k1, k2)
    nout = n // 2 if geglu else n
    if out is None:
        out = torch.empty((m, nout), dtype=x.dtype, device=x.device)
    res = residual.reshape(-1, nout) if residual is not None else None
    a = GemmArgs()
    a.a, a.a2, a.w, a.out = xm.data_ptr(), native.ptr(x2m), w.data_ptr(), out.data_ptr()
    a.bias, a.rowbias, a.residual = native.ptr(bias), native.ptr(rowbias), native.ptr(res)
    a.M, a.N, a.K1, a.K2 = m, n, k1, k2
    a.conv = 0
    a.rows_per_group = rows_per_group
    a.act, a.geglu, a.out_scale = ACT[act], int(geglu), out_scale
    a.dtype = native.dtype_code(x.dtype)
    a.block_n = block_n
    a.ldo = out.stride(0) if out.dim() == 2 else 0
    a.ldr = res.stride(0) if res is not None else 0
    a.ldrb = rowbias.stride(0) if rowbias is not None and rowbias.dim() == 2 else 0
    kc = (k1 + k2 + 63) // 64
    kind = 'geglu' if geglu else 'lin'
    key = f"{kind}:{_canon(m)}:{n}:{k1 + k2}"
    if RECORD is not None:
        RECORD.append(key)
    cfg = _tuned_pair(key, f"{kind}:{m}:{n}:{k1 + k2}") if (splitk == 0 and block_n == 0) else None
    if cfg is not None:
        a.block_n, a.splitk = cfg[0], cfg[1]
    else:
        a.splitk = splitk if splitk > 0 else auto_splitk(_canon(m), n, kc)
    if a.splitk > 1:
        a.workspace = _workspace(a.splitk * m * n * 4, x.device).data_ptr()
    native.check(lib.ab_gemm(C.byref(a), native.stream_ptr()), "ab_gemm")
    return out.reshape(*lead, nout) if out.shape[0] == m and out.dim() == 2 else out


def conv3x3(x: torch.Tensor, w_packed: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
            x2: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
            rowbias: Optional[torch.Tensor] = None, act: Optional[str] = None, out_scale: float = 1.0,
            splitk: int = 0, block_n: int = 0, halo: bool = False, sync=None, upsample: bool = False) -> torch.Tensor:
    """3x3 / pad 1 / stride 1 implicit-GEMM conv. x [B,H,W,C1] (+ x2 [B,H,W,C2]), w_packed [Cout, 9*(C1+C2)].
    ``upsample``: the conv of the nearest-2x upsampled x, [B,2H,2W,Cout], without materialising the upsampled image: a
    tile holds one parity class of the output, so every tap is a plain TMA box of x (same bytes as conv3x3 of
    ``upsample_nearest2x(x)``).
    ``sync``: (tick_ptr, site_ptr, mask, top_mask) — the kernel's TMA producer waits for those ranks' flags before its first load
    (the neighbours push their edge rows straight into this buffer's halo rows; parallel/patch.py).
    ``halo``: x (and x2) are [B, H + 2, W, C] — a row band of the image with the neighbouring bands' edge rows (or zeros
    at the image's top and bottom) in rows 0 and H+1, as a patch-parallel rank holds it; ``batch_scale`` then counts
    the bands, so the split-K choice is the whole image's."""
    lib = native.load()
    b, h, wd, c1 = x.shape
    if halo:
        h -= 2
    assert x.is_contiguous() and w_packed.is_contiguous()
    c2 = x2.shape[-1] if x2 is not None else 0
    n = w_packed.shape[0]
    assert w_packed.shape[1] == 9 * (c1 + c2)
    up = 2 if upsample else 1
    assert not (upsample and (halo or x2 is not None or sync is not None))
    out = torch.empty((b, up * h, up * wd, n), dtype=x.dtype, device=x.device)
    a = GemmArgs()
    a.a, a.a2, a.w, a.out = x.data_ptr(), native.ptr(x2), w_packed.data_ptr(), out.data_ptr()
    a.bias, a.rowbias, a.residual = native.ptr(bias), native.ptr(rowbias), native.ptr(residual)
    a.M, a.N, a.K1, a.K2 = up * up * b * h * wd, n, c1, c2
    a.conv, a.nimg, a.H, a.W = 1, b, h, wd
    a.resample = 2 if upsample else 0
    a.halo = 1 if halo else 0
    if sync is not None:
        a.sync_tick, a.sync_wait, a.sync_wait_mask, a.sync_top_mask = sync
    a.act, a.geglu, a.out_scale = ACT[act], 0, out_scale
    a.dtype = native.dtype_code(x.dtype)
    a.block_n = block_n
    a.ldrb = rowbias.stride(0) if rowbias is not None and rowbias.dim() == 2 else 0
    kc = 9 * (c1 + c2) // 64
    # (the table is keyed by the OUTPUT image: the fused upsample takes the tile / split-K of the conv it replaces)
    key = f"conv:{b}x{_canon(h)}x{wd}:{n}:{c1 + c2}" if halo else f"conv:{_canon(b)}x{up * h}x{up * wd}:{n}:{c1 + c2}"
    if RECORD is not None:
        RECORD.append(key)
    cfg = _tuned_pair(key, f"conv:{b}x{up * h}x{up * wd}:{n}:{c1 + c2}") if (splitk == 0 and block_n == 0) else None
    if cfg is not None:
        a.block_n, a.splitk = cfg[0], cfg[1]
    else:
        a.splitk = splitk if splitk > 0 else auto_splitk(_canon(a.M), n, kc)
    if a.splitk > 1:
        a.workspace = _workspace(a.splitk * a.M * n * 4, x.device).data_ptr()
    native.check(lib.ab_gemm(C.byref(a), native.stream_ptr()), "ab_gemm(conv)")
    return out


def conv3x3_strided(x: torch.Tensor, w_packed: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
                    stride: int = 2, pad_mode: int = 0, act: Optional[str] = None, splitk: int = 0,
                    block_n: int = 0) -> torch.Tensor:
    """3x3 / stride 2 conv as an implicit GEMM whose TMA boxes step two pixels per element along W and H (no gathered
    copy of the input). pad_mode 0: pad 1 all round; 1: pad right / bottom only (the VAE encoder's).
    The split-K and the K order are those of the gathered form (``conv3x3_strided_gather``), so the bytes are too."""
    if stride != 2 or x.shape[-1] % 64:
        return conv3x3_strided_gather(x, w_packed, bias, stride=stride, pad_mode=pad_mode, act=act)
    lib = native.load()
    b, h, wd, c = x.shape
    ho = (h + (2 if pad_mode == 0 else 1) - 3) // 2 + 1
    wo = (wd + (2 if pad_mode == 0 else 1) - 3) // 2 + 1
    n = w_packed.shape[0]
    assert x.is_contiguous() and w_packed.is_contiguous() and w_packed.shape[1] == 9 * c
    out = torch.empty((b, ho, wo, n), dtype=x.dtype, device=x.device)
    a = GemmArgs()
    a.a, a.w, a.out, a.bias = x.data_ptr(), w_packed.data_ptr(), out.data_ptr(), native.ptr(bias)
    a.M, a.N, a.K1, a.K2 = b * ho * wo, n, c, 0
    a.conv, a.nimg, a.H, a.W = 1, b, ho, wo
    a.resample, a.pad_lo, a.Hin, a.Win = 1, (1 if pad_mode == 0 else 0), h, wd
    a.act, a.geglu, a.out_scale = ACT[act], 0, 1.0
    a.dtype = native.dtype_code(x.dtype)
    a.block_n = block_n
    # the gathered form ran as a linear GEMM [M, 9C] x [9C, N]: same table key, so the same split-K
    m = a.M
    key = f"lin:{_canon(m)}:{n}:{9 * c}"
    if RECORD is not None:
        RECORD.append(key)
    cfg = _tuned_pair(key, f"lin:{m}:{n}:{9 * c}") if (splitk == 0 and block_n == 0) else None
    if cfg is not None:
        a.block_n, a.splitk = cfg[0], cfg[1]
    else:
        a.splitk = splitk if splitk > 0 else auto_splitk(_canon(m), n, 9 * c // 64)
    if a.splitk > 1:
        a.workspace = _workspace(a.splitk * m * n * 4, x.device).data_ptr()
    native.check(lib.ab_gemm(C.byref(a), native.stream_ptr()), "ab_gemm(conv, stride 2)")
    return out


def conv3x3_strided_gather(x: torch.Tensor, w_packed: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
                    stride: int = 2, pad_mode: int = 0, act: Optional[str] = None) -> torch.Tensor:
    """Strided 3x3 conv = gather (ab_im2col3x3) + tcgen05 GEMM. pad_mode 0: pad 1 all round; 1: right/bottom.
    The fallback of ``conv3x3_strided`` (channel counts that are not multiples of 64) and its test oracle."""
    lib = native.load()
    b, h, wd, c = x.shape
    ho = (h + (2 if pad_mode == 0 else 1) - 3) // stride + 1
    wo = (wd + (2 if pad_mode == 0 else 1) - 3) // stride + 1
    col = torch.empty((b * ho * wo, 9 * c), dtype=x.dtype, device=x.device)
    native.check(lib.ab_im2col3x3(x.data_ptr(), col.data_ptr(), b, h, wd, c, stride, pad_mode,
                                  native.dtype_code(x.dtype), native.stream_ptr()), "ab_im2col3x3")
    y = gemm(col, w_packed, bias, act=act)
    return y.reshape(b, ho, wo, -1)


def pack_temporal_conv_weight(w: torch.Tensor) -> torch.Tensor:
    """Conv3d weight [Cout, Cin, 3, 1, 1] -> [Cout, 3*Cin] with K ordered (kf, cin)."""
    co, ci = w.shape[0], w.shape[1]
    return w.reshape(co, ci, 3).permute(0, 2, 1).reshape(co, 3 * ci).contiguous()


def conv_temporal(x: torch.Tensor, w_packed: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
                  residual: Optional[torch.Tensor] = None, act: Optional[str] = None, splitk: int = 0,
                  block_n: int = 0, halo: bool = False, frame_ranks: int = 0) -> torch.Tensor:
    """Temporal Conv3d (3,1,1), pad (1,0,0), on channels-last video x [B, F, HW, C] -> [B, F, HW, Cout].
    Same implicit-GEMM kernel as conv3x3 with a 3x1 tap pattern over the (F, HW) "image".
    ``halo``: x is [B, F + 2, HW, C] — the local frames of a frame-sharded clip with the neighbours' frames (or zeros at
    the clip's ends) in rows 0 and F+1; ``frame_ranks`` is the number of ranks the frames are spread over
    (default: ``batch_scale``, which may also carry a guidance-branch split of the batch)."""
    lib = native.load()
    b, f, hw, c = x.shape
    if halo:
        f -= 2
    assert x.is_contiguous() and w_packed.is_contiguous() and w_packed.shape[1] == 3 * c
    n = w_packed.shape[0]
    out = torch.empty((b, f, hw, n), dtype=x.dtype, device=x.device)
    a = GemmArgs()
    a.a, a.w, a.out = x.data_ptr(), w_packed.data_ptr(), out.data_ptr()
    a.bias, a.residual = native.ptr(bias), native.ptr(residual)
    a.M, a.N, a.K1, a.K2 = b * f * hw, n, c, 0
    a.conv, a.nimg, a.H, a.W = 2, b, f, hw
    a.halo = 1 if halo else 0
    a.act, a.out_scale = ACT[act], 1.0
    a.dtype = native.dtype_code(x.dtype)
    if halo:
        fr = frame_ranks or int(get_batch_scale())
        key = f"tconv:{_canon(b) // fr}x{f * fr}x{hw}:{n}:{c}"
    else:
        key = f"tconv:{_canon(b)}x{f}x{hw}:{n}:{c}"
    if RECORD is not None:
        RECORD.append(key)
    cfg = tuned_config(key) if (splitk == 0 and block_n == 0) else None
    if cfg is not None:
        a.block_n, a.splitk = cfg[0], cfg[1]
    else:
        a.block_n = block_n
        a.splitk = splitk if splitk > 0 else auto_splitk(_canon(a.M), n, 3 * c // 64)
    if a.splitk > 1:
        a.workspace = _workspace(a.splitk * a.M * n * 4, x.device).data_ptr()
    native.check(lib.ab_gemm(C.byref(a), native.stream_ptr()), "ab_gemm(temporal conv)")
    return out


class DirectConvWeight:
    """Weight of a small-channel conv in the layout ab_conv_direct wants for its shape class:
    [Cout, kh, kw, Cin] when Cout <= 8 (warp-per-pixel kernel), else [kh, kw, Cin, Cout]."""

    def __init__(self, w_ohwi: torch.Tensor):
        cout, kh, kw, cin = w_ohwi.shape
        self.cout, self.cin, self.ks = cout, cin, kh
        few_out = cout <= 8 and cin % 8 == 0
        self.t = w_ohwi.contiguous() if few_out else w_ohwi.permute(1, 2, 3, 0).contiguous()


def conv_direct(x: torch.Tensor, w, bias: Optional[torch.Tensor] = None, *, ksize: Optional[int] = None,
                stride: int = 1, act: Optional[str] = None) -> torch.Tensor:
    """Small-channel conv. ``w`` is a DirectConvWeight (or a [Cout, kh, kw, Cin] tensor, packed on the fly)."""
    lib = native.load()
    if not isinstance(w, DirectConvWeight):
        w = DirectConvWeight(w)
    b, h, wd, cin = x.shape
    assert cin == w.cin and (ksize is None or ksize == w.ks)
    ks, cout = w.ks, w.cout
    pad = ks // 2
    ho, wo = (h + 2 * pad - ks) // stride + 1, (wd + 2 * pad - ks) // stride + 1
    out = torch.empty((b, ho, wo, cout), dtype=x.dtype, device=x.device)
    native.check(lib.ab_conv_direct(x.data_ptr(), w.t.data_ptr(), native.ptr(bias), out.data_ptr(), b, h, wd, cin,
                                    cout, ks, stride, ACT[act], native.dtype_code(x.dtype), native.stream_ptr()),
                 "ab_conv_direct")
    return out
