"""Flash attention vs the library (cuDNN / flash SDPA through torch) at the shapes the models run, with the SM clock
sampled after each timing. AB_ATTN_V=2 times the previous kernel.

  python scripts/bench_attn.py
"""
import json
import os
import subprocess
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
from arbius_b200 import ops  # noqa: E402


def clocks():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits", "-i", "0"],
                             capture_output=True, text=True, timeout=10).stdout.strip().split(",")
        return float(out[0]), float(out[1])
    except Exception:  # noqa: BLE001
        return 0.0, 0.0


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dt = torch.float16
    torch.manual_seed(0)
    for (b, h, l, lk, d, iters) in [(2, 8, 4096, 4096, 40, 50), (8, 8, 4096, 4096, 40, 20), (2, 16, 2304, 2304, 64, 50),
                                    (1, 5, 9216, 9216, 64, 20), (48, 5, 9216, 9216, 64, 5), (48, 10, 2304, 2304, 64, 20)]:
        q = torch.randn(b, l, h, d, device="cuda", dtype=dt)
        k = torch.randn(b, lk, h, d, device="cuda", dtype=dt)
        v = torch.randn(b, lk, h, d, device="cuda", dtype=dt)
        o = ops.flash_attn(q, k, v)
        qt, kt, vt = (z.permute(0, 2, 1, 3) for z in (q, k, v))
        ref = F.scaled_dot_product_attention(qt.float(), kt.float(), vt.float()).permute(0, 2, 1, 3) if b * h * l * lk <= 2 ** 31 else None
        err = float((o.float() - ref).abs().max()) if ref is not None else None
        t = timeit(lambda: ops.flash_attn(q, k, v), iters)
        c = clocks()
        tr = timeit(lambda: F.scaled_dot_product_attention(qt, kt, vt), iters)
        print(json.dumps(dict(op="attn", version=os.environ.get("AB_ATTN_V", "3"), shape=[b, h, l, lk, d], ms=round(t, 4),
                              tflops=round(4 * b * h * l * lk * d / t / 1e9, 1), torch_ms=round(tr, 4), max_abs_err=err,
                              sm_mhz=c[0], sm_max_mhz=c[1])), flush=True)


if __name__ == "__main__":
    main()
