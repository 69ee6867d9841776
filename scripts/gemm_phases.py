import ctypes as C, sys, torch
sys.path.insert(0, ".")
from arbius_b200 import ops
lib = ops.native.load()
lib.ab_gemm_set_debug.argtypes = [C.c_void_p]; lib.ab_gemm_set_debug.restype = None
dt = torch.float16
for (m, n, k, bn) in [(8192, 320, 320, 160), (2048, 640, 640, 96), (512, 1280, 1280, 64), (8192, 1280, 320, 160),
                      (8192, 320, 2880, 160), (8192, 8192, 8192, 256)]:
    x = torch.randn(m, k, device="cuda", dtype=dt); w = torch.randn(n, k, device="cuda", dtype=dt); b = torch.randn(n, device="cuda", dtype=dt)
    for _ in range(2): ops.gemm(x, w, b, block_n=bn, splitk=1)
    torch.cuda.synchronize()
    dbg = torch.zeros(8 * 65536, dtype=torch.int64, device="cuda")
    lib.ab_gemm_set_debug(dbg.data_ptr())
    ops.gemm(x, w, b, block_n=bn, splitk=1)
    torch.cuda.synchronize()
    lib.ab_gemm_set_debug(None)
    nct = ((m + 127) // 128) * ((n + bn - 1) // bn)
    d = dbg[: 8 * nct].view(nct, 8).float()
    print((m, n, k, bn), "ctas", nct, "median cycles [setup, first_full, mma_issued, acc_ready, epi_done, dealloc]:",
          [int(v) for v in d.median(0).values[:6].tolist()], "max", [int(v) for v in d.max(0).values[:6].tolist()])
