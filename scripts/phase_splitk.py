import ctypes as C, sys, torch
sys.path.insert(0, ".")
sys.path.insert(0, "scripts")
from arbius_b200 import ops
from arbius_b200.ops import native
import gemm_phases as gp
lib = native.load()
lib.ab_gemm_set_cluster_splitk.argtypes = [C.c_int]
torch.manual_seed(0)
for (b, h, w, c, n) in [(2, 48, 48, 768, 768), (2, 12, 12, 1536, 1536)]:
    x = torch.randn(b, h, w, c, device="cuda").half()
    wt = ops.pack_conv_weight((torch.randn(n, c, 3, 3, device="cuda") * 0.02).half())
    bias = torch.randn(n, device="cuda").half()
    for bn in (128, 256):
        for sk in (2, 4, 6):
            for mode in (0, 1):
                lib.ab_gemm_set_cluster_splitk(mode)
                gp.run(f"conv {b}x{h}x{w} {c}->{n} bn{bn} sk{sk} {'cluster' if mode else 'workspace'}",
                       lambda: ops.conv3x3(x, wt, bias, block_n=bn, splitk=sk))
