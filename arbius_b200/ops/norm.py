"""GroupNorm(+SiLU) / LayerNorm wrappers (csrc/norm.cu)."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import native
from .native import ACT, GroupNormArgs

_gn_ws: dict = {}


def _gn_workspace(n: int, groups: int, device) -> torch.Tensor:
    lib = native.load()
    need = int(lib.ab_groupnorm_partials_bytes(n, groups))
    key = (device.index if device.index is not None else torch.cuda.current_device(), n, groups)
    ws = _gn_ws.get(key)
    if ws is None:
        # zero-filled once: the kernel's block counters reset themselves after every call
        ws = torch.zeros(need, dtype=torch.uint8, device=device)
        _gn_ws[key] = ws
    return ws


def groupnorm(x: torch.Tensor, gamma: Optional[torch.Tensor], beta: Optional[torch.Tensor], groups: int = 32,
              eps: float = 1e-5, act: Optional[str] = None, *, x2: Optional[torch.Tensor] = None,
              scale: Optional[torch.Tensor] = None, shift: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [N, ..., C1] channels-last (+ x2 [N, ..., C2] virtual concat) -> [N, ..., C1+C2]."""
    lib = native.load()
    n, c1 = x.shape[0], x.shape[-1]
    hw = x.numel() // (n * c1)
    c2 = x2.shape[-1] if x2 is not None else 0
    assert x.is_contiguous() and (x2 is None or x2.is_contiguous())
    y = torch.empty(x.shape[:-1] + (c1 + c2,), dtype=x.dtype, device=x.device)
    a = GroupNormArgs()
    a.x, a.x2, a.y = x.data_ptr(), native.ptr(x2), y.data_ptr()
    a.gamma, a.beta = native.ptr(gamma), native.ptr(beta)
    a.partials = _gn_workspace(n, groups, x.device).data_ptr()
    a.n, a.hw, a.c1, a.c2, a.groups = n, hw, c1, c2, groups
    a.eps, a.act = eps, ACT[act]
    a.dtype = native.dtype_code(x.dtype)
    a.scale, a.shift = native.ptr(scale), native.ptr(shift)
    if scale is not None and scale.dim() == 2:
        assert shift is None or shift.stride(0) == scale.stride(0)
        a.ss_stride = scale.stride(0)              # scale/shift may be column slices of one fused projection
    native.check(lib.ab_groupnorm(C.byref(a), native.stream_ptr()), "ab_groupnorm")
    return y


def layernorm(x: torch.Tensor, gamma: Optional[torch.Tensor], beta: Optional[torch.Tensor],
              eps: float = 1e-5) -> torch.Tensor:
    lib = native.load()
    cols = x.shape[-1]
    rows = x.numel() // cols
    assert x.is_contiguous()
    y = torch.empty_like(x)
    native.check(lib.ab_layernorm(x.data_ptr(), native.ptr(gamma), native.ptr(beta), y.data_ptr(), rows, cols, eps,
                                  native.dtype_code(x.dtype), native.stream_ptr()), "ab_layernorm")
    return y
