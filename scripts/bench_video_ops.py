"""ZeroScope top-level memory-bound ops: GroupNorm(+SiLU), temporal attention, LayerNorm — time and achieved GB/s."""
import sys, torch
sys.path.insert(0, ".")
from arbius_b200 import ops
from arbius_b200.ops.temporal import temporal_attn
dt = torch.float16


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for (n, hw, c) in [(48, 9216, 320), (48, 2304, 640), (48, 9216, 640)]:
    x = torch.randn(n, hw, c, device="cuda", dtype=dt)
    g = torch.randn(c, device="cuda", dtype=dt)
    ms = timeit(lambda: ops.groupnorm(x, g, g, 32, 1e-5, "silu"))
    print(f"groupnorm+silu {n}x{hw}x{c}: {ms * 1e3:.0f} us  {2 * x.numel() * 2 / ms / 1e6:.0f} GB/s (read once + write)", flush=True)
    ms = timeit(lambda: ops.layernorm(x.view(-1, c), g, g))
    print(f"layernorm      {n * hw}x{c}: {ms * 1e3:.0f} us  {2 * x.numel() * 2 / ms / 1e6:.0f} GB/s", flush=True)
for (b, f, hw, heads, d) in [(2, 24, 9216, 5, 64), (2, 24, 2304, 10, 64)]:
    c = heads * d
    qkv = torch.randn(b, f, hw, 3 * c, device="cuda", dtype=dt)
    q, k, v = (qkv[..., i * c:(i + 1) * c] for i in range(3))
    ms = timeit(lambda: temporal_attn(q, k, v, heads))
    print(f"temporal_attn  B{b} F{f} HW{hw} {heads}x{d}: {ms * 1e3:.0f} us  {4 * b * f * hw * c * 2 / ms / 1e6:.0f} GB/s (q,k,v read + out)", flush=True)
