"""Video-scale implicit-GEMM convs (ZeroScope top levels): time and TFLOP/s; run with AB_GEMM_MT2=0/1 to compare."""
import sys, torch
sys.path.insert(0, ".")
from arbius_b200 import ops
dt = torch.float16
for (n, h, w, ci, co) in [(48, 72, 128, 320, 320), (48, 72, 128, 640, 320), (48, 36, 64, 640, 640), (48, 18, 32, 1280, 1280)]:
    x = torch.randn(n, h, w, ci, device="cuda", dtype=dt)
    wt = torch.randn(co, 9 * ci, device="cuda", dtype=dt) * 0.01
    b = torch.randn(co, device="cuda", dtype=dt)
    for _ in range(3):
        ops.conv3x3(x, wt, b)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ops.conv3x3(x, wt, b)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    fl = 2.0 * n * h * w * ci * 9 * co
    print(f"conv3x3 {n}x{h}x{w} {ci}->{co}: {ms:.3f} ms  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
for (n, h, w, ci, co) in [(48, 18, 32, 1280, 1280), (48, 36, 64, 640, 640), (48, 72, 128, 320, 512)]:
    x = torch.randn(n, h, w, ci, device="cuda", dtype=dt)
    wt = torch.randn(co, 9 * ci, device="cuda", dtype=dt) * 0.01
    b = torch.randn(co, device="cuda", dtype=dt)
    for _ in range(3):
        ops.conv3x3(x, wt, b, block_n=256)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ops.conv3x3(x, wt, b, block_n=256)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    fl = 2.0 * n * h * w * ci * 9 * co
    print(f"conv3x3 BN=256 {n}x{h}x{w} {ci}->{co}: {ms:.3f} ms  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
