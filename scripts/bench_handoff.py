"""NVLink hand-off latency between two B200s (torchrun, 2 ranks): flag only, 16-byte flag-in-data entry, and
`bytes` of peer stores + system fence + flag (a halo push). Prints one-way latencies in microseconds."""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arbius_b200.ops import native  # noqa: E402
from arbius_b200.parallel import init_distributed  # noqa: E402
from arbius_b200.parallel.symm import PeerHeap  # noqa: E402


def main():
    world = init_distributed()
    assert world == 2
    rank = dist.get_rank()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    heap = PeerHeap(64 << 20)
    flags, foff = heap.alloc((64,), torch.int32)
    buf, boff = heap.alloc((4 << 20,), torch.uint8)
    lib = native.load()
    vp = C.c_void_p
    lib.ab_sync_pingpong.restype = C.c_int
    lib.ab_sync_pingpong.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint, vp, vp]
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    base = 0
    for mode, nbytes, label in [(0, 0, "flag only (st.release.sys -> ld.acquire.sys)"), (1, 0, "16-byte flag-in-data entry"),
                                (2, 4096, "4 KB peer stores + system fence + flag"),
                                (2, 147456, "144 KB peer stores + system fence + flag (one 96x384 halo row x 2 images)")]:
        iters = 2000 if mode < 2 else 500
        for rep in range(2):
            dist.barrier()
            torch.cuda.synchronize()
            e0.record()
            native.check(lib.ab_sync_pingpong(flags.data_ptr(), heap.peer_ptr(1 - rank, foff), heap.peer_ptr(1 - rank, boff), rank,
                                              iters, mode, nbytes, base, out.data_ptr(), native.stream_ptr()), "pingpong")
            e1.record()
            torch.cuda.synchronize()
            base += iters + 10
        cyc = out[0].item()
        us = e0.elapsed_time(e1) * 1e3 / iters / 2
        if rank == 0:
            print(f"RESULT handoff {label}: {us:.2f} us one way (CUDA events; {cyc / iters / 2:.0f} SM cycles)", flush=True)
    heap.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
