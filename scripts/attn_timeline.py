"""Timeline of one CTA of the v3 flash attention kernel (clock64 stamps through ab_attn_set_debug): when the MMA
thread sees each tile's P, when the softmax warps get S and finish their exponentials. Cycles relative to the first stamp.

  python scripts/attn_timeline.py
"""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from arbius_b200 import ops  # noqa: E402
from arbius_b200.ops import native  # noqa: E402


def main():
    lib = native.load()
    lib.ab_attn_set_debug.argtypes = [C.c_void_p]
    lib.ab_attn_set_debug.restype = None
    b, l, h, d = 2, 9216, 5, 64
    q = torch.randn(b, l, h, d, device="cuda", dtype=torch.float16)
    k = torch.randn_like(q)
    v = torch.randn_like(q)
    for _ in range(3):
        ops.flash_attn(q, k, v)
    buf = torch.zeros(4096, dtype=torch.int64, device="cuda")
    lib.ab_attn_set_debug(buf.data_ptr())
    ops.flash_attn(q, k, v)
    torch.cuda.synchronize()
    lib.ab_attn_set_debug(None)
    t = buf.cpu()
    mma = t[:128].view(32, 4)
    sa = t[1024:1088].view(32, 2)
    sb = t[2048:2112].view(32, 2)
    t0 = int(sa[0, 0])
    print("tile  | MMA: P_A seen  A issued  P_B seen  B issued | softmax A: S ready  done (compute) | softmax B: S ready  done (compute)")
    for j in range(4, 20):
        m = [int(x) - t0 for x in mma[j]]
        a = [int(x) - t0 for x in sa[j]]
        bb = [int(x) - t0 for x in sb[j]]
        print(f"{j:4d}  | {m[0]:9d} {m[1]:9d} {m[2]:9d} {m[3]:9d} | {a[0]:9d} {a[1]:9d} ({a[1] - a[0]:5d}) | {bb[0]:9d} {bb[1]:9d} ({bb[1] - bb[0]:5d})")
    print("cycles per KV tile (A, tiles 4..31):", (int(sa[31, 0]) - int(sa[4, 0])) / 27)


if __name__ == "__main__":
    main()
