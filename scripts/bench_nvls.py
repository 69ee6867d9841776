"""NVLS broadcast vs per-peer stores (torchrun, 2/4/8 ranks): every rank sends `bytes` to all the others, as a
patch-parallel attention layer does with its band's K/V. Checks the received bytes, prints device-timed microseconds
(max over ranks). If the driver offers no multicast on this box, says so and exits 0."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arbius_b200.parallel import init_distributed  # noqa: E402
from arbius_b200.parallel import patch as pp  # noqa: E402
from arbius_b200.parallel.nvls import MulticastBuffer  # noqa: E402
from arbius_b200.parallel.symm import PeerHeap  # noqa: E402


def main():
    world = init_distributed()
    rank = dist.get_rank()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if not MulticastBuffer.supported():
        if rank == 0:
            print("RESULT nvls unsupported: CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 0 on this box", flush=True)
        dist.destroy_process_group()
        return
    for nbytes in (64 << 10, 1769472, 8 << 20):              # 1769472 = one 48x48 band's K|V at P=8 (288 x 2 x 768 x 2 x 2)
        try:
            mc = MulticastBuffer(world * nbytes)
        except RuntimeError as e:
            if rank == 0:
                print(f"RESULT nvls setup failed: {e}", flush=True)
            break
        src = torch.full((nbytes // 2,), float(rank + 1), dtype=torch.float16, device="cuda")
        dist.barrier()
        mc.broadcast(src, rank * nbytes)
        torch.cuda.synchronize()
        dist.barrier()
        got = mc.local[:world * nbytes].view(torch.float16).view(world, -1)      # (the buffer is rounded up to the granularity)
        ok = all(bool((got[r] == float(r + 1)).all()) for r in range(world))
        t = torch.tensor([0.0 if ok else 1.0], device="cuda")
        dist.all_reduce(t)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 50

        def timeit(fn):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            dist.barrier()
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            v = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], device="cuda")
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            return v.item()
        us_mc = timeit(lambda: mc.broadcast(src, rank * nbytes))
        heap = PeerHeap(world * nbytes + (1 << 20))
        ctx = pp.PatchContext(heap, torch.float16) if world in (2, 4, 8) else None
        buf, off = heap.alloc((world, nbytes), torch.uint8)
        segs = [(heap.peer_ptr(r, off) + rank * nbytes, src.data_ptr(), nbytes, nbytes, nbytes, 1) for r in range(world) if r != rank]
        us_p2p = timeit(lambda: ctx.copy(segs)) if ctx is not None else float("nan")
        if rank == 0:
            print(f"RESULT nvls world={world} bytes={nbytes} received_ok={t.item() == 0} multimem_st_us={us_mc:.2f} "
                  f"per_peer_stores_us={us_p2p:.2f} outbound_GBps_multimem={nbytes / us_mc / 1e3:.1f} "
                  f"outbound_GBps_per_peer={(world - 1) * nbytes / us_p2p / 1e3:.1f}", flush=True)
        heap.close()
        mc.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
