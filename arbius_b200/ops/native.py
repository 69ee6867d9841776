"""ctypes binding to ``libarbius_b200.so`` (built in-tree by ``csrc/build.sh``).

The library exposes a plain C ABI (``include/arbius_b200.h``); tensors cross the
boundary as raw device pointers plus the current CUDA stream, so nothing here
depends on torch's C++ ABI and a rebuild takes seconds. On a GPU box a missing
library is a hard error: there is no silent PyTorch fallback for the native ops.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
import subprocess
import threading
from pathlib import Path
from typing import Optional

_ROOT = Path(__file__).resolve().parents[2]
_LIB_PATH = Path(__file__).resolve().parents[1] / "_native" / "libarbius_b200.so"
_lib: Optional[C.CDLL] = None

F16, BF16 = 0, 1
ACT = {None: 0, "none": 0, "silu": 1, "gelu": 2, "quick_gelu": 3, "gelu_tanh": 4, "sigmoid": 5, "tanh": 6,
       "relu": 7, "hardswish": 8, "hardsigmoid": 9, "relu6": 10}


class GemmArgs(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("a2", C.c_void_p), ("w", C.c_void_p), ("out", C.c_void_p),
        ("bias", C.c_void_p), ("rowbias", C.c_void_p), ("residual", C.c_void_p),
        ("workspace", C.c_void_p),
        ("M", C.c_int), ("N", C.c_int), ("K1", C.c_int), ("K2", C.c_int),
        ("conv", C.c_int), ("nimg", C.c_int), ("H", C.c_int), ("W", C.c_int),
        ("rows_per_group", C.c_int),
        ("act", C.c_int), ("geglu", C.c_int),
        ("out_scale", C.c_float),
        ("dtype", C.c_int),
        ("splitk", C.c_int),
        ("block_n", C.c_int),
        ("ldo", C.c_int),
        ("ldr", C.c_int),
        ("ldrb", C.c_int),
        ("halo", C.c_int),
        ("sync_tick", C.c_void_p), ("sync_wait", C.c_void_p), ("sync_wait_mask", C.c_uint), ("sync_top_mask", C.c_uint),
        ("resample", C.c_int), ("pad_lo", C.c_int), ("Hin", C.c_int), ("Win", C.c_int),
    ]


class GroupNormArgs(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("x2", C.c_void_p), ("y", C.c_void_p),
        ("gamma", C.c_void_p), ("beta", C.c_void_p),
        ("partials", C.c_void_p),
        ("n", C.c_int), ("hw", C.c_int), ("c1", C.c_int), ("c2", C.c_int), ("groups", C.c_int),
        ("eps", C.c_float),
        ("act", C.c_int),
        ("dtype", C.c_int),
        ("scale", C.c_void_p), ("shift", C.c_void_p),
        ("ss_stride", C.c_int),
    ]


class AttnArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("out", C.c_void_p),
        ("batch", C.c_int), ("heads", C.c_int), ("lq", C.c_int), ("lk", C.c_int), ("head_dim", C.c_int),
        ("q_bs", C.c_longlong), ("q_ls", C.c_longlong), ("k_bs", C.c_longlong), ("k_ls", C.c_longlong),
        ("v_bs", C.c_longlong), ("v_ls", C.c_longlong), ("o_bs", C.c_longlong), ("o_ls", C.c_longlong),
        ("scale", C.c_float),
        ("causal", C.c_int),
        ("dtype", C.c_int),
    ]


_CAPTURE_LOCK = threading.RLock()


@contextlib.contextmanager
def graph_capture(graph):
    """``torch.cuda.graph`` for pipelines that may be driven from several threads (one predictor per GPU in one
    process): captures are serialised process-wide (concurrent ``capture_begin`` calls fail with
    cudaErrorIllegalState) and use thread-local capture mode so the other threads' replays and allocations neither
    break the capture nor are broken by it. Replays stay concurrent."""
    import torch
    with _CAPTURE_LOCK:
        # An explicit stream on the CURRENT device: torch.cuda.graph's default capture stream is created once per
        # process on whichever device was current then, and entering it silently switches devices.
        # capture_begin/capture_end are called directly: the torch.cuda.graph context also empties the caching
        # allocator on entry, and the cudaFree calls behind that wait for EVERY device of the process to go idle —
        # while another thread's GPU may be spinning in a peer barrier for this thread (a deadlock until the barrier's
        # timeout; seen with the in-process pair predictor).
        stream = torch.cuda.Stream(device=torch.cuda.current_device())
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            graph.capture_begin(capture_error_mode="thread_local")
            try:
                yield
            finally:
                graph.capture_end()
        torch.cuda.current_stream().wait_stream(stream)


def lib_path() -> Path:
    return _LIB_PATH


def build(verbose: bool = False) -> Path:
    """Compile every .cu under csrc/ for sm_100a into the in-tree shared library."""
    res = subprocess.run(["bash", str(_ROOT / "csrc" / "build.sh")], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"native build failed:\n{res.stdout}\n{res.stderr}")
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def _sig(fn, restype, argtypes):
    fn.restype = restype
    fn.argtypes = argtypes


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        if os.environ.get("ARBIUS_B200_AUTOBUILD", "1") == "1":
            build()
        else:
            raise RuntimeError(f"{_LIB_PATH} is missing; run csrc/build.sh")
    lib = C.CDLL(str(_LIB_PATH))
    vp, i, f, ll, ull = C.c_void_p, C.c_int, C.c_float, C.c_longlong, C.c_ulonglong
    _sig(lib.ab_last_error, C.c_char_p, [])
    _sig(lib.ab_version, i, [])
    _sig(lib.ab_launch_count, ull, [])
    _sig(lib.ab_gemm, i, [C.POINTER(GemmArgs), vp])
    _sig(lib.ab_gemm_workspace_bytes, C.c_size_t, [C.POINTER(GemmArgs)])
    _sig(lib.ab_gemm_set_2sm, None, [i, i])          # two-SM kernel on / off and its row threshold (tests, A/B timing)
    _sig(lib.ab_gemm_set_debug, None, [vp])          # profiling hooks: scripts/gemm_phases.py, scripts/attn_timeline.py
    _sig(lib.ab_attn_set_debug, None, [vp])
    _sig(lib.ab_conv_direct, i, [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, vp])
    _sig(lib.ab_groupnorm, i, [C.POINTER(GroupNormArgs), vp])
    _sig(lib.ab_groupnorm_partials_bytes, C.c_size_t, [i, i])
    _sig(lib.ab_softmax_rows, i, [vp, vp, i, i, f, i, vp])
    _sig(lib.ab_transpose2d, i, [vp, vp, i, i, i, i, vp])
    _sig(lib.ab_embedding, i, [vp, vp, vp, vp, i, i, i, i, vp])
    _sig(lib.ab_layernorm, i, [vp, vp, vp, vp, i, i, f, i, vp])
    if hasattr(lib, "ab_flash_attn"):
        _sig(lib.ab_flash_attn, i, [C.POINTER(AttnArgs), vp])
    _sig(lib.ab_philox_normal, i, [vp, ll, ull, ull, ll, vp])
    _sig(lib.ab_cfg_step, i, [vp, vp, vp, vp, vp, vp, vp, vp, ll, f, f, f, f, f, i, vp])
    _sig(lib.ab_lincomb, i, [vp, vp, f, C.POINTER(vp), C.POINTER(f), i, ll, i, vp])
    _sig(lib.ab_finalize_u8, i, [vp, vp, ll, i, vp])
    _sig(lib.ab_upsample_nearest2x, i, [vp, vp, i, i, i, i, i, vp])
    _sig(lib.ab_resize_bilinear, i, [vp, vp, i, i, i, i, i, i, i, i, vp])
    _sig(lib.ab_im2col3x3, i, [vp, vp, i, i, i, i, i, i, i, vp])
    _sig(lib.ab_binary, i, [vp, vp, vp, ll, i, f, i, i, vp])
    _sig(lib.ab_timestep_embedding, i, [vp, vp, i, i, i, f, f, i, vp])
    _sig(lib.ab_matting_composite, i, [vp, vp, vp, ll, i, i, vp])
    _lib = lib
    return lib


def check(rc: int, what: str = "native call") -> None:
    if rc != 0:
        msg = load().ab_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed (rc={rc}): {msg}")


def launch_count() -> int:
    return int(load().ab_launch_count())


def dtype_code(dt) -> int:
    import torch
    if dt == torch.float16:
        return F16
    if dt == torch.bfloat16:
        return BF16
    raise TypeError(f"native kernels take fp16/bf16 tensors, got {dt}")


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> Optional[int]:
    return None if t is None else t.data_ptr()
