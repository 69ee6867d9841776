"""The two-SM GEMM (csrc/gemm_2sm.cu) against the one-SM kernel (AB_GEMM_2SM=0) on the large projections of ZeroScope XL
and of the four-task SD replay: microseconds per call in a CUDA graph, and a hash of the output bytes (must be equal).

  python scripts/bench_2sm.py; AB_GEMM_2SM=0 python scripts/bench_2sm.py
"""
import hashlib
import json
import os
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "scripts")
from arbius_b200 import ops  # noqa: E402
from bench_persist import REPS, time_graph  # noqa: E402

torch.manual_seed(0)
dt = torch.float16
for kind, m, n, k, bn in [("lin", 442368, 320, 320, 160), ("geglu", 442368, 2560, 320, 256), ("geglu", 110592, 5120, 640, 256),
                          ("geglu", 27648, 10240, 1280, 256), ("lin", 442368, 320, 1280, 160), ("lin", 110592, 640, 640, 160),
                          ("lin", 27648, 1280, 1280, 160), ("lin", 32768, 320, 320, 160), ("geglu", 32768, 2560, 320, 256),
                          ("lin", 32768, 1280, 320, 256), ("lin", 32768, 320, 1280, 160), ("lin", 16384, 640, 640, 160)]:
    x = torch.randn(m, k, device="cuda", dtype=dt)
    w = torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5
    b = torch.randn(n, device="cuda", dtype=dt)

    def fn():
        for _ in range(REPS):
            ops.gemm(x, w, b, geglu=(kind == "geglu"), block_n=bn, splitk=1)
    us = time_graph(fn)
    y = ops.gemm(x, w, b, geglu=(kind == "geglu"), block_n=bn, splitk=1)
    h = hashlib.sha1(y.cpu().numpy().tobytes()).hexdigest()[:10]
    print(json.dumps(dict(two_sm=os.environ.get("AB_GEMM_2SM", "1"), key=f"{kind}:{m}:{n}:{k}", us=round(us, 1),
                          tflops=round(2.0 * m * n * k / us / 1e6, 1), sha1=h)), flush=True)
    del x, w, b, y
