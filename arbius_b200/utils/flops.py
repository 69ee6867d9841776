"""FLOP accounting from the engine's own call log (``ops.gemm.RECORD``): every tensor-core call appends a shape key
(``lin:M:N:K``, ``geglu:M:N:K``, ``conv:BxHxW:N:C``, ``tconv:BxFxHW:N:C``, ``attn:B:H:Lq:Lk:D``), so the work of a task
is counted from the calls the engine really makes — the numerator of the roofline fractions bench.py reports against
the measured cuBLAS peak (MEASURED_PEAKS.json). Canonical-batch keys carry the canonical M, which is what a lone task
computes; elementwise, normalisation and small direct convs are not counted (they are bandwidth, not tensor work)."""
from __future__ import annotations

from typing import Iterable


def flops_of_key(key: str) -> float:
    p = key.split(":")
    kind = p[0]
    if kind in ("lin", "geglu"):
        m, n, k = (int(v) for v in p[1:4])
        return 2.0 * m * n * k
    if kind in ("conv", "tconv"):
        b, h, w = (int(v) for v in p[1].split("x"))
        n, c = int(p[2]), int(p[3])
        return 2.0 * b * h * w * n * c * (9 if kind == "conv" else 3)
    if kind == "attn":
        b, h, lq, lk, d = (int(v) for v in p[1:6])
        return 4.0 * b * h * lq * lk * d
    return 0.0


def total_flops(keys: Iterable[str]) -> float:
    return sum(flops_of_key(k) for k in keys)
