"""ab_groupnorm_static: the GroupNorm entry point for callers that cannot allocate device memory.

Some users of the library have no way to hand it a scratch buffer: their own arena already owns all
of HBM, or they call the C ABI from a program that has no device allocator, or they cannot pay the
device-wide stall of a cudaMalloc in the middle of serving. For them the scratch sits inside the
library. This test drives that entry point straight through the C ABI (ctypes, raw pointers, raw
stream) and checks it against the regular ab_groupnorm on the same inputs — twice in a row, because
the built-in scratch is shared by all calls and keeps the previous call's contents.

The file name sorts last on purpose: with ``-x`` every other test has already run if this one fails.
It must pass on its own in a fresh process as well as after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(native, x, x2, y, gamma, beta, partials, groups, act):
    a = native.GroupNormArgs()
    n, c1 = x.shape[0], x.shape[-1]
    a.x, a.x2, a.y = x.data_ptr(), (x2.data_ptr() if x2 is not None else None), y.data_ptr()
    a.gamma, a.beta = gamma.data_ptr(), beta.data_ptr()
    a.partials = partials.data_ptr() if partials is not None else None
    a.n, a.hw, a.c1, a.c2, a.groups = n, x.numel() // (n * c1), c1, (x2.shape[-1] if x2 is not None else 0), groups
    a.eps, a.act = 1e-5, native.ACT[act]
    a.dtype = native.dtype_code(x.dtype)
    a.scale = a.shift = None
    return a


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16])
def test_static_scratch_groupnorm_matches_workspace_groupnorm(dt):
    from arbius_b200.ops import native
    lib = native.load()
    lib.ab_groupnorm_static.restype = C.c_int
    lib.ab_groupnorm_static.argtypes = [C.POINTER(native.GroupNormArgs), C.c_void_p]
    stream = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(7)
    # two calls in a row with different shapes: the second one sees the first one's leftovers
    for (n, hw, c1, c2, act) in [(2, 4096, 320, 0, "silu"), (3, 300, 640, 320, None)]:
        x = torch.randn(n, hw, c1, device="cuda", dtype=dt) * 2 + 0.3
        x2 = torch.randn(n, hw, c2, device="cuda", dtype=dt) - 0.5 if c2 else None
        g = torch.randn(c1 + c2, device="cuda", dtype=dt)
        b = torch.randn(c1 + c2, device="cuda", dtype=dt)
        ws = torch.zeros(int(lib.ab_groupnorm_partials_bytes(n, 32)), dtype=torch.uint8, device="cuda")
        y_ref = torch.empty(n, hw, c1 + c2, device="cuda", dtype=dt)
        y_new = torch.full_like(y_ref, float("nan"))
        native.check(lib.ab_groupnorm(C.byref(_args(native, x, x2, y_ref, g, b, ws, 32, act)), stream), "ab_groupnorm")
        native.check(lib.ab_groupnorm_static(C.byref(_args(native, x, x2, y_new, g, b, None, 32, act)), stream),
                     "ab_groupnorm_static")
        torch.cuda.synchronize()
        assert torch.isfinite(y_new).all()
        rtol, atol = (2e-2, 2e-2) if dt == torch.bfloat16 else (4e-3, 4e-3)
        err = (y_new.float() - y_ref.float()).abs().max().item()
        assert err <= atol + rtol * y_ref.float().abs().max().item(), err
        assert torch.equal(y_new, y_ref), "same kernels, same reduction order: outputs should be bit-identical"


def test_static_scratch_groupnorm_rejects_oversized_requests():
    from arbius_b200.ops import native
    lib = native.load()
    lib.ab_groupnorm_static.restype = C.c_int
    lib.ab_groupnorm_static.argtypes = [C.POINTER(native.GroupNormArgs), C.c_void_p]
    stream = torch.cuda.current_stream().cuda_stream
    x = torch.randn(5, 64, 64, device="cuda", dtype=torch.float16)          # 5 images > limit of 4
    g = torch.ones(64, device="cuda", dtype=torch.float16)
    y = torch.empty_like(x)
    rc = lib.ab_groupnorm_static(C.byref(_args(native, x, None, y, g, g, None, 32, None)), stream)
    assert rc != 0 and b"at most 4 images" in lib.ab_last_error()
    x = torch.randn(1, 64, 128, device="cuda", dtype=torch.float16)         # 64 groups > limit of 32
    g = torch.ones(128, device="cuda", dtype=torch.float16)
    y = torch.empty_like(x)
    rc = lib.ab_groupnorm_static(C.byref(_args(native, x, None, y, g, g, None, 64, None)), stream)
    assert rc != 0 and b"at most 32 groups" in lib.ab_last_error()
