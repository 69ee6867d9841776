"""Wrappers for csrc/elementwise2.cu: pooling, SpatialNorm modulation, ConvGRU update, depthwise conv, SE helpers."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import native
from .native import ACT


def _lib():
    lib = native.load()
    if not hasattr(lib, "_extra_sigs"):
        vp, i, ll = C.c_void_p, C.c_int, C.c_longlong
        for name, args in (("ab_avgpool2x", [vp, vp, i, i, i, i, i, vp]),
                           ("ab_spatial_modulate", [vp, vp, vp, vp, i, i, i, i, i, i, i, i, vp]),
                           ("ab_gru_update", [vp, vp, vp, vp, ll, i, vp]),
                           ("ab_dwconv", [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, vp]),
                           ("ab_global_avgpool", [vp, vp, i, i, i, i, vp]),
                           ("ab_channel_scale", [vp, vp, vp, i, i, i, i, vp])):
            fn = getattr(lib, name)
            fn.restype = i
            fn.argtypes = args
        lib._extra_sigs = True
    return lib


def avgpool2x(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    n, h, w, c = x.shape
    y = out if out is not None else torch.empty((n, h // 2, w // 2, c), dtype=x.dtype, device=x.device)
    assert x.is_contiguous() and y.is_contiguous() and y.shape == (n, h // 2, w // 2, c)
    native.check(_lib().ab_avgpool2x(x.data_ptr(), y.data_ptr(), n, h, w, c, native.dtype_code(x.dtype),
                                     native.stream_ptr()), "ab_avgpool2x")
    return y


def spatial_modulate(x: torch.Tensor, ymap: torch.Tensor, bmap: torch.Tensor, act: Optional[str] = None) -> torch.Tensor:
    n, h, w, c = x.shape
    lh, lw = ymap.shape[1], ymap.shape[2]
    out = torch.empty_like(x)
    native.check(_lib().ab_spatial_modulate(x.data_ptr(), ymap.data_ptr(), bmap.data_ptr(), out.data_ptr(), n, h, w, c,
                                            lh, lw, ACT[act], native.dtype_code(x.dtype), native.stream_ptr()),
                 "ab_spatial_modulate")
    return out


def gru_update(z: torch.Tensor, cand: torch.Tensor, hprev: torch.Tensor) -> torch.Tensor:
    out = torch.empty_like(hprev)
    native.check(_lib().ab_gru_update(z.data_ptr(), cand.data_ptr(), hprev.data_ptr(), out.data_ptr(), hprev.numel(),
                                      native.dtype_code(hprev.dtype), native.stream_ptr()), "ab_gru_update")
    return out


def dwconv(x: torch.Tensor, w_kkc: torch.Tensor, bias: Optional[torch.Tensor], ksize: int, stride: int = 1,
           dilation: int = 1, act: Optional[str] = None) -> torch.Tensor:
    """x [N,H,W,C], w_kkc [K*K, C] -> [N,Ho,Wo,C]."""
    n, h, w, c = x.shape
    pad = (ksize // 2) * dilation
    ho = (h + 2 * pad - dilation * (ksize - 1) - 1) // stride + 1
    wo = (w + 2 * pad - dilation * (ksize - 1) - 1) // stride + 1
    y = torch.empty((n, ho, wo, c), dtype=x.dtype, device=x.device)
    native.check(_lib().ab_dwconv(x.data_ptr(), w_kkc.data_ptr(), native.ptr(bias), y.data_ptr(), n, h, w, c, ksize,
                                  stride, dilation, ACT[act], native.dtype_code(x.dtype), native.stream_ptr()), "ab_dwconv")
    return y


def global_avgpool(x: torch.Tensor) -> torch.Tensor:
    """[N, ..., C] -> [N, C]"""
    n, c = x.shape[0], x.shape[-1]
    hw = x.numel() // (n * c)
    y = torch.empty((n, c), dtype=x.dtype, device=x.device)
    native.check(_lib().ab_global_avgpool(x.data_ptr(), y.data_ptr(), n, hw, c, native.dtype_code(x.dtype),
                                          native.stream_ptr()), "ab_global_avgpool")
    return y


def channel_scale(x: torch.Tensor, s: torch.Tensor) -> torch.Tensor:
    n, c = x.shape[0], x.shape[-1]
    hw = x.numel() // (n * c)
    y = torch.empty_like(x)
    native.check(_lib().ab_channel_scale(x.data_ptr(), s.data_ptr(), y.data_ptr(), n, hw, c, native.dtype_code(x.dtype),
                                         native.stream_ptr()), "ab_channel_scale")
    return y
