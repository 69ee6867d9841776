"""Per-stage timing the way every number in this repository is taken (the reference has no timers at
all): CUDA events on the launching stream, read after a synchronize, and — when several ranks take part — the MAX
over ranks, never a wall clock. On a box without a GPU the same interface falls back to ``perf_counter`` so the
control-plane tests can use it.

    t = StageTimer()
    with t.stage("text_encoder"): ...
    with t.stage("unet"): ...
    print(t.summary())            # {"text_encoder": ms, "unet": ms}, max over ranks if torch.distributed is up
"""
from __future__ import annotations

import contextlib
import math
import time
from typing import Dict, List, Optional


class StageTimer:
    def __init__(self, device: Optional[str] = None):
        import torch
        self._cuda = torch.cuda.is_available() and (device is None or str(device).startswith("cuda"))
        self._events: Dict[str, List[tuple]] = {}
        self._host: Dict[str, float] = {}

    @contextlib.contextmanager
    def stage(self, name: str):
        if self._cuda:
            import torch
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            try:
                yield
            finally:
                b.record()
                self._events.setdefault(name, []).append((a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self._host[name] = self._host.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def summary(self, reduce_ranks: bool = True) -> Dict[str, float]:
        """Milliseconds per stage (sum over its occurrences); max over ranks when a process group is up."""
        out = dict(self._host)
        if self._cuda:
            import torch
            torch.cuda.synchronize()
            for name, evs in self._events.items():
                out[name] = out.get(name, 0.0) + sum(a.elapsed_time(b) for a, b in evs)
        if reduce_ranks:
            import torch
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                names = sorted(out)
                t = torch.tensor([out[n] for n in names], dtype=torch.float64,
                                 device="cuda" if self._cuda and dist.get_backend() == "nccl" else "cpu")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                out = {n: float(v) for n, v in zip(names, t.tolist())}
        return out

    def reset(self) -> None:
        self._events.clear()
        self._host.clear()


class LatencyHistogram:
    """Log-spaced latency buckets with exact count/sum and interpolated quantiles (p50 is a headline metric); renders
    as Prometheus text for the miner's ``GET /metrics``."""

    def __init__(self, lo_ms: float = 1.0, hi_ms: float = 600_000.0, per_decade: int = 10):
        n = int(math.ceil(math.log10(hi_ms / lo_ms) * per_decade))
        self.bounds = [lo_ms * 10 ** (i / per_decade) for i in range(n + 1)]
        self.counts = [0] * (len(self.bounds) + 1)
        self.n, self.total = 0, 0.0

    def observe(self, ms: float) -> None:
        self.n += 1
        self.total += ms
        lo, hi = 0, len(self.bounds)
        while lo < hi:
            mid = (lo + hi) // 2
            if ms <= self.bounds[mid]:
                hi = mid
            else:
                lo = mid + 1
        self.counts[lo] += 1

    def quantile(self, q: float) -> float:
        if self.n == 0:
            return 0.0
        target, acc = q * self.n, 0
        for i, c in enumerate(self.counts):
            if acc + c >= target and c > 0:
                lo = self.bounds[i - 1] if i > 0 else 0.0
                hi = self.bounds[i] if i < len(self.bounds) else self.bounds[-1]
                return lo + (hi - lo) * (target - acc) / c
            acc += c
        return self.bounds[-1]

    def prometheus(self, name: str, labels: str = "") -> str:
        lab = (labels + ",") if labels else ""
        lines, acc = [], 0
        for b, c in zip(self.bounds, self.counts):
            acc += c
            lines.append(f'{name}_bucket{{{lab}le="{b:.6g}"}} {acc}')
        lines.append(f'{name}_bucket{{{lab}le="+Inf"}} {self.n}')
        lines.append(f"{name}_sum{{{labels}}} {self.total:.6g}")
        lines.append(f"{name}_count{{{labels}}} {self.n}")
        return "\n".join(lines)
