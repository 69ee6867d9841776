"""Elementwise / data-movement wrappers (csrc/elementwise.cu) and the CPU twin of the Philox RNG."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import native
from .native import ACT

# ---------------------------------------------------------------------------------------------
# Philox4x32-10, counter = global element index / 4 — CPU twin of philox_normal_kernel.
# Integer stream is bit-identical to the GPU; the Box-Muller floats agree to fp32 rounding of
# log/sincos (the CID only ever depends on the GPU path; the twin is the oracle for
# shard-invariance and for CPU-only tests). SURVEY §7.3 item 2.
# ---------------------------------------------------------------------------------------------
_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85


def philox_u32(seed: int, stream_id: int, first_counter: int, count: int) -> np.ndarray:
    """uint32 [count, 4]: Philox output blocks for counters first_counter .. first_counter+count-1."""
    q = np.arange(first_counter, first_counter + count, dtype=np.uint64)
    c0 = (q & np.uint64(0xFFFFFFFF)).astype(np.uint64)
    c1 = (q >> np.uint64(32)).astype(np.uint64)
    c2 = np.full(count, stream_id & 0xFFFFFFFF, dtype=np.uint64)
    c3 = np.full(count, (stream_id >> 32) & 0xFFFFFFFF, dtype=np.uint64)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(_M0) * c0
        p1 = np.uint64(_M1) * c2
        n0 = ((p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0)) & mask
        n1 = p1 & mask
        n2 = ((p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1)) & mask
        n3 = p0 & mask
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return np.stack([c0, c1, c2, c3], axis=1).astype(np.uint32)


def philox_normal_cpu(n: int, seed: int, stream_id: int = 0, offset: int = 0) -> torch.Tensor:
    first_q, last_q = offset >> 2, (offset + n - 1) >> 2
    u = philox_u32(seed, stream_id, first_q, last_q - first_q + 1)
    f = ((u >> 8).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)
    z = np.empty((u.shape[0], 4), dtype=np.float32)
    for h in range(2):
        r = np.sqrt(np.float32(-2.0) * np.log(f[:, 2 * h])).astype(np.float32)
        ang = (np.float32(2.0) * f[:, 2 * h + 1]).astype(np.float32) * np.float32(np.pi)
        z[:, 2 * h] = r * np.cos(ang).astype(np.float32)
        z[:, 2 * h + 1] = r * np.sin(ang).astype(np.float32)
    flat = z.reshape(-1)
    start = offset - 4 * first_q
    return torch.from_numpy(flat[start:start + n].copy())


def philox_normal(shape, seed: int, stream_id: int = 0, offset: int = 0, device="cuda") -> torch.Tensor:
    """fp32 N(0,1) tensor whose element i depends only on (seed, stream_id, offset + i)."""
    n = int(np.prod(shape))
    dev = torch.device(device)
    if dev.type != "cuda":
        return philox_normal_cpu(n, seed, stream_id, offset).reshape(shape)
    lib = native.load()
    out = torch.empty(n, dtype=torch.float32, device=dev)
    native.check(lib.ab_philox_normal(out.data_ptr(), n, seed, stream_id, offset, native.stream_ptr()),
                 "ab_philox_normal")
    return out.reshape(shape)


def cfg_step(eps_c: torch.Tensor, eps_u: Optional[torch.Tensor], x: Optional[torch.Tensor], *, g: float = 1.0,
             ca: float = 0.0, cb: float = 1.0, aux: Optional[torch.Tensor] = None, cc: float = 0.0,
             noise: Optional[torch.Tensor] = None, cn: float = 0.0, x_out: Optional[torch.Tensor] = None,
             xm_out: Optional[torch.Tensor] = None, eps_out: Optional[torch.Tensor] = None):
    """eps = eps_u + g (eps_c - eps_u); x_out = ca x + cb eps + cc aux + cn noise (fp32)."""
    lib = native.load()
    n = eps_c.numel()
    native.check(lib.ab_cfg_step(eps_c.data_ptr(), native.ptr(eps_u), native.ptr(x), native.ptr(aux),
                                 native.ptr(noise), native.ptr(x_out), native.ptr(xm_out), native.ptr(eps_out), n,
                                 g, ca, cb, cc, cn, native.dtype_code(eps_c.dtype), native.stream_ptr()),
                 "ab_cfg_step")
    return x_out


def lincomb(dst: Optional[torch.Tensor], terms, *, dst_m: Optional[torch.Tensor] = None, m_scale: float = 1.0):
    """dst = sum coef*src over fp32 tensors (scheduler plan op). terms: [(coef, tensor), ...], <= 6."""
    import ctypes as C
    lib = native.load()
    k = len(terms)
    srcs = (C.c_void_p * k)(*[t.data_ptr() for _, t in terms])
    coefs = (C.c_float * k)(*[float(c) for c, _ in terms])
    n = terms[0][1].numel()
    code = native.dtype_code(dst_m.dtype) if dst_m is not None else 0
    native.check(lib.ab_lincomb(native.ptr(dst), native.ptr(dst_m), m_scale, srcs, coefs, k, n, code,
                                native.stream_ptr()), "ab_lincomb")
    return dst


def finalize_u8(x: torch.Tensor) -> torch.Tensor:
    lib = native.load()
    out = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    native.check(lib.ab_finalize_u8(x.data_ptr(), out.data_ptr(), x.numel(), native.dtype_code(x.dtype),
                                    native.stream_ptr()), "ab_finalize_u8")
    return out


def upsample_nearest2x(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    lib = native.load()
    n, h, w, c = x.shape
    y = out if out is not None else torch.empty((n, 2 * h, 2 * w, c), dtype=x.dtype, device=x.device)
    assert x.is_contiguous() and y.is_contiguous() and y.shape == (n, 2 * h, 2 * w, c)
    native.check(lib.ab_upsample_nearest2x(x.data_ptr(), y.data_ptr(), n, h, w, c, native.dtype_code(x.dtype),
                                           native.stream_ptr()), "ab_upsample_nearest2x")
    return y


def resize_bilinear(x: torch.Tensor, oh: int, ow: int, align_corners: bool = False) -> torch.Tensor:
    lib = native.load()
    n, h, w, c = x.shape
    y = torch.empty((n, oh, ow, c), dtype=x.dtype, device=x.device)
    native.check(lib.ab_resize_bilinear(x.data_ptr(), y.data_ptr(), n, h, w, c, oh, ow, int(align_corners),
                                        native.dtype_code(x.dtype), native.stream_ptr()), "ab_resize_bilinear")
    return y


def binary(a: torch.Tensor, b: Optional[torch.Tensor], op: str = "add", s: float = 1.0,
           act: Optional[str] = None) -> torch.Tensor:
    lib = native.load()
    y = torch.empty_like(a)
    code = {"add": 0, "mul": 1, "axpy": 2}[op]
    native.check(lib.ab_binary(a.data_ptr(), native.ptr(b), y.data_ptr(), a.numel(), code, s, ACT[act],
                               native.dtype_code(a.dtype), native.stream_ptr()), "ab_binary")
    return y


def timestep_embedding(t: torch.Tensor, dim: int, dtype: torch.dtype, flip_sin_to_cos: bool = True,
                       freq_shift: float = 0.0, max_period: float = 10000.0) -> torch.Tensor:
    lib = native.load()
    t = t.to(torch.float32).contiguous()
    out = torch.empty((t.numel(), dim), dtype=dtype, device=t.device)
    native.check(lib.ab_timestep_embedding(t.data_ptr(), out.data_ptr(), t.numel(), dim, int(flip_sin_to_cos),
                                           freq_shift, max_period, native.dtype_code(dtype), native.stream_ptr()),
                 "ab_timestep_embedding")
    return out


def matting_composite(fgr: torch.Tensor, pha: torch.Tensor, mode: str = "green-screen") -> torch.Tensor:
    lib = native.load()
    code = {"green-screen": 0, "": 0, "alpha-mask": 1, "foreground-mask": 2}[mode]
    npix = pha.numel()
    out = torch.empty(fgr.shape, dtype=torch.uint8, device=fgr.device)
    native.check(lib.ab_matting_composite(fgr.data_ptr(), pha.data_ptr(), out.data_ptr(), npix, code,
                                          native.dtype_code(fgr.dtype), native.stream_ptr()), "ab_matting_composite")
    return out
