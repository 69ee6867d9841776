"""Short-K GEMMs with and without the M-persistent launch (AB_GEMM_PERSIST=0 disables it): microseconds per call in a
CUDA graph and a hash of the output bytes (must not depend on the launch shape).

  python scripts/bench_persist.py; AB_GEMM_PERSIST=0 python scripts/bench_persist.py
"""
import hashlib
import json
import os
import sys

import torch

sys.path.insert(0, ".")
from arbius_b200 import ops  # noqa: E402

REPS = 8


def time_graph(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e3 / REPS


def main():
    torch.manual_seed(0)
    dt = torch.float16
    for kind, m, n, k, bn in [("lin", 442368, 320, 320, 160), ("lin", 442368, 960, 320, 160), ("geglu", 442368, 2560, 320, 160),
                              ("lin", 110592, 640, 640, 160), ("geglu", 110592, 5120, 640, 128), ("lin", 32768, 320, 320, 160),
                              ("geglu", 32768, 2560, 320, 160), ("lin", 32768, 1280, 320, 160), ("lin", 8192, 320, 320, 160),
                              ("lin", 442368, 320, 1280, 160), ("lin", 1000, 320, 320, 64), ("lin", 40000, 96, 128, 96)]:
        x = torch.randn(m, k, device="cuda", dtype=dt)
        w = torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5
        b = torch.randn(n, device="cuda", dtype=dt)
        res = torch.randn(m, n // 2 if kind == "geglu" else n, device="cuda", dtype=dt)
        y = ops.gemm(x, w, b, geglu=(kind == "geglu"), block_n=bn, splitk=1, residual=None if kind == "geglu" else res,
                     act=None if kind == "geglu" else "silu")
        torch.cuda.synchronize()
        h = hashlib.sha1(y.cpu().numpy().tobytes()).hexdigest()[:12]

        def fn():
            for _ in range(REPS):
                ops.gemm(x, w, b, geglu=(kind == "geglu"), block_n=bn, splitk=1)
        us = time_graph(fn)
        fl = 2.0 * m * n * k
        print(json.dumps(dict(persist=os.environ.get("AB_GEMM_PERSIST", "1"), key=f"{kind}:{m}:{n}:{k}", bn=bn, us=round(us, 1),
                              tflops=round(fl / us / 1e6, 1), sha1=h)), flush=True)
        del x, w, b, res, y
    b_, h_, w_, c, n = 8, 64, 64, 320, 320      # a conv with a long K loop: never persistent, unchanged
    x = torch.randn(b_, h_, w_, c, device="cuda", dtype=dt)
    wt = ops.pack_conv_weight((torch.randn(n, c, 3, 3, device="cuda") * 0.02).half())
    y = ops.conv3x3(x, wt, None)
    print(json.dumps(dict(key="conv:8x64x64:320:320", sha1=hashlib.sha1(y.cpu().numpy().tobytes()).hexdigest()[:12])))
    b_, h_, w_, c, n = 48, 72, 128, 64, 64      # a conv with a short K loop (9 steps): persistent
    x = torch.randn(b_, h_, w_, c, device="cuda", dtype=dt)
    wt = ops.pack_conv_weight((torch.randn(n, c, 3, 3, device="cuda") * 0.05).half())
    bias = torch.randn(n, device="cuda", dtype=dt)
    y = ops.conv3x3(x, wt, bias, act="silu")
    print(json.dumps(dict(key="conv:48x72x128:64:64", sha1=hashlib.sha1(y.cpu().numpy().tobytes()).hexdigest()[:12])))


if __name__ == "__main__":
    main()
