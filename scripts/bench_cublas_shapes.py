"""cuBLAS (torch.nn.functional.linear) on the short-K projection shapes, graph-timed like scripts/bench_persist.py."""
import json
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
sys.path.insert(0, "scripts")
from bench_persist import time_graph, REPS
for m, n, k in [(442368, 320, 320), (442368, 960, 320), (442368, 2560, 320), (110592, 640, 640), (110592, 5120, 640),
                (32768, 320, 320), (32768, 2560, 320), (442368, 320, 1280)]:
    x = torch.randn(m, k, device="cuda", dtype=torch.float16)
    w = torch.randn(n, k, device="cuda", dtype=torch.float16)
    b = torch.randn(n, device="cuda", dtype=torch.float16)
    def fn():
        for _ in range(REPS):
            F.linear(x, w, b)
    us = time_graph(fn)
    print(json.dumps(dict(lib="cublas", key=f"lin:{m}:{n}:{k}", us=round(us, 1), tflops=round(2.0 * m * n * k / us / 1e6, 1))), flush=True)
    del x, w, b
