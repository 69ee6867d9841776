"""Kernel micro-benchmarks (CUDA events, L2 flushed between iterations). Writes JSON lines to stdout."""
import json
import subprocess
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
from arbius_b200 import ops  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def sm_mhz():
    """Instantaneous SM clock and throttle reasons (sampled right after a timed loop, while the GPU is still warm)."""
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.active", "--format=csv,noheader,nounits",
                              "-i", "0"], capture_output=True, text=True, timeout=10).stdout.strip().split(",")
        return dict(sm_mhz=float(out[0]), sm_max_mhz=float(out[1]), reasons=out[2].strip())
    except Exception:  # noqa: BLE001
        return {}


def timeit(fn, iters=20, warmup=5):
    """Median device time of ONE call, L2 flushed before each. The call is replayed from a CUDA graph (both arms), so the
    host-side launch path — ctypes here, the dispatcher there — is not part of a 10 us kernel's time."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    run = fn
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        run = g.replay
        run()
        torch.cuda.synchronize()
    except Exception:  # noqa: BLE001 - an op that cannot be captured is timed eagerly
        torch.cuda.synchronize()
        run = fn
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    dt = torch.float16
    rows = []
    for (m, n, k) in [(8192, 320, 320), (8192, 2560, 320), (8192, 320, 1280), (2048, 640, 640), (2048, 5120, 640),
                      (512, 1280, 1280), (512, 10240, 1280), (128, 1280, 1280), (8192, 8192, 8192),
                      (32768, 320, 320), (442368, 320, 320), (442368, 320, 1280), (110592, 640, 640)]:
        x = torch.randn(m, k, device="cuda", dtype=dt)
        w = torch.randn(n, k, device="cuda", dtype=dt)
        b = torch.randn(n, device="cuda", dtype=dt)
        t = timeit(lambda: ops.gemm(x, w, b))
        tr = timeit(lambda: F.linear(x, w, b))
        rows.append(dict(op="gemm", m=m, n=n, k=k, ms=t, tflops=2 * m * n * k / t / 1e9, torch_ms=tr))
    for (b_, h, w_, ci, co) in [(2, 64, 64, 320, 320), (2, 64, 64, 640, 320), (2, 32, 32, 640, 640),
                                (2, 32, 32, 1280, 640), (2, 16, 16, 1280, 1280), (2, 16, 16, 2560, 1280),
                                (2, 8, 8, 1280, 1280), (1, 256, 256, 256, 256), (1, 512, 512, 128, 128)]:
        x = torch.randn(b_, h, w_, ci, device="cuda", dtype=dt)
        wt = torch.randn(co, ci, 3, 3, device="cuda", dtype=dt)
        wp = ops.pack_conv_weight(wt)
        bias = torch.randn(co, device="cuda", dtype=dt)
        t = timeit(lambda: ops.conv3x3(x, wp, bias))
        xc = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        wc = wt.contiguous(memory_format=torch.channels_last)
        tr = timeit(lambda: F.conv2d(xc, wc, bias, padding=1))
        fl = 2 * b_ * h * w_ * ci * co * 9
        rows.append(dict(op="conv3x3", shape=[b_, h, w_, ci, co], ms=t, tflops=fl / t / 1e9, torch_ms=tr))
    for (b_, h, l, lk, d) in [(2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80), (2, 8, 256, 256, 160),
                              (2, 8, 4096, 77, 40), (1, 5, 9216, 9216, 64), (48, 5, 9216, 9216, 64)]:
        q = torch.randn(b_, l, h, d, device="cuda", dtype=dt)
        k = torch.randn(b_, lk, h, d, device="cuda", dtype=dt)
        v = torch.randn(b_, lk, h, d, device="cuda", dtype=dt)
        t = timeit(lambda: ops.flash_attn(q, k, v), iters=10)
        qt, kt, vt = (z.permute(0, 2, 1, 3) for z in (q, k, v))
        tr = timeit(lambda: F.scaled_dot_product_attention(qt, kt, vt), iters=10)
        fl = 4 * b_ * h * l * lk * d
        rows.append(dict(op="attn", shape=[b_, h, l, lk, d], ms=t, tflops=fl / t / 1e9, torch_ms=tr))
    for (n, hw, c) in [(2, 4096, 320), (2, 1024, 640), (2, 256, 1280), (1, 262144, 128)]:
        x = torch.randn(n, hw, c, device="cuda", dtype=dt)
        g = torch.randn(c, device="cuda", dtype=dt)
        t = timeit(lambda: ops.groupnorm(x, g, g, 32, 1e-5, "silu"))
        xc = x.permute(0, 2, 1).reshape(n, c, hw, 1).contiguous(memory_format=torch.channels_last)
        tr = timeit(lambda: F.silu(F.group_norm(xc, 32, g, g, 1e-5)))
        rows.append(dict(op="gn_silu", shape=[n, hw, c], ms=t, gbs=2 * x.numel() * 2 / t / 1e6, torch_ms=tr))
    clocks = sm_mhz()
    for r in rows:
        r["clocks_at_end"] = clocks
        print(json.dumps(r))


if __name__ == "__main__":
    main()
