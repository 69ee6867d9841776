"""Tiny driver for ncu captures: a few GEMM / conv / attention / GroupNorm calls at UNet shapes."""
import sys
import torch
sys.path.insert(0, ".")
from arbius_b200 import ops
what = sys.argv[1] if len(sys.argv) > 1 else "gemm"
dt = torch.float16
torch.manual_seed(0)
if what == "gemm":
    for (m, n, k, bn) in [(8192, 320, 320, 160), (8192, 1280, 320, 160), (2048, 640, 640, 96)]:
        x = torch.randn(m, k, device="cuda", dtype=dt); w = torch.randn(n, k, device="cuda", dtype=dt); b = torch.randn(n, device="cuda", dtype=dt)
        for _ in range(3):
            ops.gemm(x, w, b, block_n=bn, splitk=1)
elif what == "gemmbig":    # ZeroScope's short-K projections: M = 442368 rows, K = 320 (five K steps per tile), M-persistent launch
    for (m, n, k, bn) in [(442368, 320, 320, 160)]:
        x = torch.randn(m, k, device="cuda", dtype=dt); w = torch.randn(n, k, device="cuda", dtype=dt); b = torch.randn(n, device="cuda", dtype=dt)
        for _ in range(3):
            ops.gemm(x, w, b, block_n=bn, splitk=1)
elif what == "conv":
    for (b_, h, w_, ci, co, bn) in [(2, 64, 64, 320, 320, 160), (2, 32, 32, 640, 640, 160), (1, 256, 256, 256, 256, 160)]:
        x = torch.randn(b_, h, w_, ci, device="cuda", dtype=dt); wt = torch.randn(co, 9 * ci, device="cuda", dtype=dt); b = torch.randn(co, device="cuda", dtype=dt)
        for _ in range(3):
            ops.conv3x3(x, wt, b, block_n=bn, splitk=1)
elif what == "convb":      # the UNet's 32x32 level at four tasks per replay (batch 8): the largest share of the batched task
    for (b_, h, w_, ci, co, bn, sk) in [(8, 32, 32, 1280, 1280, 160, 2), (8, 64, 64, 640, 640, 160, 1)]:
        x = torch.randn(b_, h, w_, ci, device="cuda", dtype=dt); wt = torch.randn(co, 9 * ci, device="cuda", dtype=dt) / (9 * ci) ** 0.5; b = torch.randn(co, device="cuda", dtype=dt)
        for _ in range(3):
            ops.conv3x3(x, wt, b, block_n=bn, splitk=sk)
elif what == "attn":
    shape = (2, 9216, 5, 64) if (len(sys.argv) > 2 and sys.argv[2] == "video") else (2, 4096, 8, 40)
    q = torch.randn(*shape, device="cuda", dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)
    for _ in range(3):
        ops.flash_attn(q, k, v)
elif what == "gn":
    x = torch.randn(2, 4096, 320, device="cuda", dtype=dt); g = torch.randn(320, device="cuda", dtype=dt)
    for _ in range(3):
        ops.groupnorm(x, g, g, 32, 1e-5, "silu")
torch.cuda.synchronize()
print("ok")
if what == "tattn":
    from arbius_b200.ops.temporal import temporal_attn
    b, f, hw, heads, d = 2, 24, 9216, 5, 64
    c = heads * d
    qkv = torch.randn(b, f, hw, 3 * c, device="cuda", dtype=dt)
    q, k, v = (qkv[..., i * c:(i + 1) * c] for i in range(3))
    for _ in range(3):
        temporal_attn(q, k, v, heads)
    torch.cuda.synchronize()
    print("ok")
if what == "tattn_ms":     # the frame-parallel form on one GPU: two sources of 12 frames each, rank 0's pixel band
    from arbius_b200.ops.temporal import temporal_attn_ptrs
    b, f, hw, heads, d, world = 2, 24, 9216, 5, 64, 2
    c = heads * d
    srcs = [torch.randn(b, f // world, hw, 3 * c, device="cuda", dtype=dt) for _ in range(world)]
    outs = [torch.zeros(b, f // world, hw, c, device="cuda", dtype=dt) for _ in range(world)]
    qp = [s.data_ptr() for s in srcs]
    px = 128 // f
    per = ((hw + px - 1) // px + world - 1) // world
    for _ in range(3):
        temporal_attn_ptrs(qp, [p + c * 2 for p in qp], [p + 2 * c * 2 for p in qp], [o.data_ptr() for o in outs],
                           b, f, hw, heads, d, 3 * c, c, dt, pix_begin=0, pix_count=min(hw, per * px))
    torch.cuda.synchronize()
    print("ok")
