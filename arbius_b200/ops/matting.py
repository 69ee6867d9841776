"""Wrappers for csrc/matting.cu: small-channel direct conv (virtual concat / gates / ConvGRU epilogue), explicit-step
bilinear resize, ceil-mode pooling, strided copies and the deep guided filter."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import native
from .native import ACT


class ConvGenericArgs(C.Structure):
    _fields_ = [
        ("src", C.c_void_p * 3), ("gate", C.c_void_p * 3),
        ("c", C.c_int * 3), ("ld", C.c_int * 3), ("gate_ld", C.c_int * 3),
        ("nsrc", C.c_int),
        ("w", C.c_void_p), ("bias", C.c_void_p), ("in_scale", C.c_void_p), ("in_shift", C.c_void_p),
        ("out", C.c_void_p), ("out_ld", C.c_int),
        ("gru_z", C.c_void_p), ("gru_h", C.c_void_p), ("gru_z_ld", C.c_int), ("gru_h_ld", C.c_int),
        ("n", C.c_int), ("h", C.c_int), ("w_in", C.c_int), ("ks", C.c_int), ("stride", C.c_int),
        ("dilation", C.c_int), ("cout", C.c_int), ("act", C.c_int), ("dtype", C.c_int),
    ]


def _lib():
    lib = native.load()
    if not hasattr(lib, "_matting_sigs"):
        vp, i, ll, f = C.c_void_p, C.c_int, C.c_longlong, C.c_float
        for name, args in (("ab_conv_generic", [C.POINTER(ConvGenericArgs), vp]),
                           ("ab_resize_bilinear_ex", [vp, i, vp, i, i, i, i, i, i, i, f, f, i, i, vp]),
                           ("ab_avgpool2x_ceil", [vp, vp, i, i, i, i, i, vp]),
                           ("ab_copy_cols", [vp, i, vp, i, vp, i, ll, i, i, vp]),
                           ("ab_guided_stats", [vp, vp, vp, vp, vp, i, i, i, i, vp]),
                           ("ab_guided_apply", [vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, i, i, i, vp])):
            fn = getattr(lib, name)
            fn.restype = i
            fn.argtypes = args
        lib._matting_sigs = True
    return lib


def pack_generic_weight(w: torch.Tensor, dtype, device) -> torch.Tensor:
    """torch conv weight [Cout, Cin, KH, KW] -> [KH*KW*Cin, round_up(Cout, 16)]"""
    cout, cin, kh, kw = w.shape
    cp = (cout + 15) // 16 * 16
    out = torch.zeros(kh * kw * cin, cp, dtype=dtype, device=device)
    out[:, :cout] = w.permute(2, 3, 1, 0).reshape(kh * kw * cin, cout).to(device=device, dtype=dtype)
    return out.contiguous()


def conv_generic(srcs: Sequence[torch.Tensor], w_packed: torch.Tensor, bias: Optional[torch.Tensor], cout: int, ks: int, *,
                 stride: int = 1, dilation: int = 1, act: Optional[str] = None,
                 gates: Optional[Sequence[Optional[torch.Tensor]]] = None,
                 in_affine: Optional[tuple] = None, gru: Optional[tuple] = None,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """srcs: NHWC tensors (possibly channel-sliced views of wider NHWC tensors) -> [N, Ho, Wo, cout]."""
    x0 = srcs[0]
    n, h, w = x0.shape[0], x0.shape[1], x0.shape[2]
    pad = (ks // 2) * dilation
    ho = (h + 2 * pad - dilation * (ks - 1) - 1) // stride + 1
    wo = (w + 2 * pad - dilation * (ks - 1) - 1) // stride + 1
    if out is None:
        out = torch.empty((n, ho, wo, cout), dtype=x0.dtype, device=x0.device)
    a = ConvGenericArgs()
    a.nsrc = len(srcs)
    ctot = 0
    for k, s in enumerate(srcs):
        assert s.shape[:3] == x0.shape[:3] and s.stride(-1) == 1 and _pixel_major(s), "source must be an NHWC (sliced) view"
        a.src[k], a.c[k], a.ld[k] = s.data_ptr(), s.shape[-1], s.stride(2)
        g = gates[k] if gates is not None else None
        if g is not None:
            assert g.shape == s.shape and g.stride(-1) == 1 and _pixel_major(g)
            a.gate[k], a.gate_ld[k] = g.data_ptr(), g.stride(2)
        ctot += s.shape[-1]
    assert w_packed.shape[0] == ks * ks * ctot and w_packed.shape[1] == (cout + 15) // 16 * 16, (w_packed.shape, ctot, cout)
    a.w, a.bias = w_packed.data_ptr(), native.ptr(bias)
    if in_affine is not None:
        a.in_scale, a.in_shift = in_affine[0].data_ptr(), in_affine[1].data_ptr()
    assert out.stride(-1) == 1 and _pixel_major(out) and out.shape[-1] == cout
    a.out, a.out_ld = out.data_ptr(), out.stride(2)
    if gru is not None:
        z, hprev = gru
        assert z.shape[-1] == cout and hprev.shape[-1] == cout and _pixel_major(z) and _pixel_major(hprev)
        a.gru_z, a.gru_h, a.gru_z_ld, a.gru_h_ld = z.data_ptr(), hprev.data_ptr(), z.stride(2), hprev.stride(2)
    a.n, a.h, a.w_in, a.ks, a.stride, a.dilation, a.cout = n, h, w, ks, stride, dilation, cout
    a.act, a.dtype = ACT[act], native.dtype_code(x0.dtype)
    native.check(_lib().ab_conv_generic(C.byref(a), native.stream_ptr()), "ab_conv_generic")
    return out


def _pixel_major(t: torch.Tensor) -> bool:
    """[N,H,W,C] view whose pixels are laid out contiguously with one row stride (C may be a slice)."""
    ld = t.stride(2)
    return t.dim() == 4 and t.stride(1) == ld * t.shape[2] and t.stride(0) == ld * t.shape[2] * t.shape[1]


def resize_bilinear_ex(x: torch.Tensor, oh: int, ow: int, step_h: float, step_w: float, dtype=None,
                       out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [N,H,W,C] (model dtype or uint8 -> /255). src = (dst+0.5)*step-0.5 like align_corners=False."""
    n, h, w, c = x.shape
    u8 = x.dtype == torch.uint8
    odt = dtype if dtype is not None else x.dtype
    assert not (u8 and dtype is None), "uint8 input needs an output dtype"
    if out is None:
        out = torch.empty((n, oh, ow, c), dtype=odt, device=x.device)
    assert _pixel_major(x) and _pixel_major(out) and x.stride(-1) == 1 and out.stride(-1) == 1
    native.check(_lib().ab_resize_bilinear_ex(x.data_ptr(), x.stride(2), out.data_ptr(), out.stride(2), n, h, w, c, oh, ow,
                                              float(step_h), float(step_w), int(u8), native.dtype_code(out.dtype),
                                              native.stream_ptr()), "ab_resize_bilinear_ex")
    return out


def avgpool2x_ceil(x: torch.Tensor) -> torch.Tensor:
    n, h, w, c = x.shape
    assert x.is_contiguous()
    y = torch.empty((n, (h + 1) // 2, (w + 1) // 2, c), dtype=x.dtype, device=x.device)
    native.check(_lib().ab_avgpool2x_ceil(x.data_ptr(), y.data_ptr(), n, h, w, c, native.dtype_code(x.dtype),
                                          native.stream_ptr()), "ab_avgpool2x_ceil")
    return y


def copy_cols(src: torch.Tensor, dst1: torch.Tensor, dst2: Optional[torch.Tensor] = None) -> None:
    """copy an NHWC (sliced) view to one or two NHWC (sliced) views of the same shape"""
    assert _pixel_major(src) and _pixel_major(dst1) and src.shape == dst1.shape
    rows, cols = src.shape[0] * src.shape[1] * src.shape[2], src.shape[3]
    d2, l2 = (dst2.data_ptr(), dst2.stride(2)) if dst2 is not None else (None, 0)
    native.check(_lib().ab_copy_cols(src.data_ptr(), src.stride(2), dst1.data_ptr(), dst1.stride(2), d2, l2, rows, cols,
                                     native.dtype_code(src.dtype), native.stream_ptr()), "ab_copy_cols")


def guided_stats(base_src: torch.Tensor, proj: torch.Tensor):
    n, h, w, _ = base_src.shape
    assert base_src.is_contiguous() and proj.is_contiguous() and base_src.shape[-1] == 3 and proj.shape[-1] == 4
    stats = torch.empty((n, h, w, 8), dtype=proj.dtype, device=proj.device)
    mx = torch.empty((n, h, w, 4), dtype=proj.dtype, device=proj.device)
    my = torch.empty_like(mx)
    native.check(_lib().ab_guided_stats(base_src.data_ptr(), proj.data_ptr(), stats.data_ptr(), mx.data_ptr(), my.data_ptr(),
                                        n, h, w, native.dtype_code(proj.dtype), native.stream_ptr()), "ab_guided_stats")
    return stats, mx, my


MODES = {"green-screen": 0, "alpha-mask": 1, "foreground-mask": 2}


def guided_apply(src_u8: torch.Tensor, a_or_proj: torch.Tensor, mean_x: Optional[torch.Tensor], mean_y: Optional[torch.Tensor],
                 *, want_fgr_pha: bool = False, mode: Optional[int] = None, out_u8: Optional[torch.Tensor] = None):
    """src_u8 [N,H,W,3] -> (fgr [N,H,W,3], pha [N,H,W,1]) and/or composited uint8 frame."""
    n, H, W, _ = src_u8.shape
    refine = mean_x is not None
    h, w = a_or_proj.shape[1], a_or_proj.shape[2]
    assert src_u8.dtype == torch.uint8 and src_u8.is_contiguous() and a_or_proj.is_contiguous()
    fgr = pha = None
    if want_fgr_pha:
        fgr = torch.empty((n, H, W, 3), dtype=a_or_proj.dtype, device=src_u8.device)
        pha = torch.empty((n, H, W, 1), dtype=a_or_proj.dtype, device=src_u8.device)
    if mode is not None and out_u8 is None:
        out_u8 = torch.empty((n, H, W, 3), dtype=torch.uint8, device=src_u8.device)
    native.check(_lib().ab_guided_apply(src_u8.data_ptr(), a_or_proj.data_ptr(), native.ptr(mean_x), native.ptr(mean_y),
                                        native.ptr(fgr), native.ptr(pha), native.ptr(out_u8), n, H, W, h, w, int(refine),
                                        int(mode or 0), native.dtype_code(a_or_proj.dtype), native.stream_ptr()),
                 "ab_guided_apply")
    return fgr, pha, out_u8
