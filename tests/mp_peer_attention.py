"""torchrun worker: patch-parallel attention over the peer heap must equal the single-GPU kernel bit for bit."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arbius_b200 import ops  # noqa: E402
from arbius_b200.parallel import init_distributed  # noqa: E402
from arbius_b200.parallel.patch_attention import PatchAttention  # noqa: E402
from arbius_b200.parallel.symm import PeerHeap  # noqa: E402


def main():
    world = init_distributed()
    rank = dist.get_rank()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    heap = PeerHeap(64 << 20)
    ok = True
    for (b, l, h, d) in [(2, 4096, 8, 40), (1, 2048, 5, 64), (2, 1024, 4, 32)]:
        torch.manual_seed(1234)                      # every rank builds the same full tensors
        q = torch.randn(b, l, h, d, device="cuda", dtype=torch.float16)
        k = torch.randn(b, l, h, d, device="cuda", dtype=torch.float16)
        v = torch.randn(b, l, h, d, device="cuda", dtype=torch.float16)
        ref = ops.flash_attn(q, k, v)                # single-GPU kernel on the full problem
        ll = l // world
        sl = slice(rank * ll, (rank + 1) * ll)
        heap.reset()
        pa = PatchAttention(heap, b, ll, h, d)
        out = pa(q[:, sl].contiguous(), k[:, sl].contiguous(), v[:, sl].contiguous())
        torch.cuda.synchronize()
        same = torch.equal(out, ref[:, sl])
        err = (out.float() - ref[:, sl].float()).abs().max().item()
        print(f"rank {rank} shape {(b, l, h, d)} bit-identical={same} maxerr={err:.3e}", flush=True)
        ok = ok and same
    # timing: fused (peer TMA) vs NCCL all-gather + local attention, device time, max over ranks
    b, l, h, d = 2, 4096, 8, 40
    ll = l // world
    q = torch.randn(b, ll, h, d, device="cuda", dtype=torch.float16)
    k = torch.randn(b, ll, h, d, device="cuda", dtype=torch.float16)
    v = torch.randn(b, ll, h, d, device="cuda", dtype=torch.float16)
    heap.reset()
    pa = PatchAttention(heap, b, ll, h, d)

    def nccl_path():
        ks = [torch.empty_like(k) for _ in range(world)]
        vs = [torch.empty_like(v) for _ in range(world)]
        dist.all_gather(ks, k)
        dist.all_gather(vs, v)
        return ops.flash_attn(q, torch.cat(ks, 1), torch.cat(vs, 1))

    def timeit(fn, iters=20):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()
    t_fused = timeit(lambda: pa(q, k, v))
    t_nccl = timeit(nccl_path)
    if rank == 0:
        print(f"RESULT world={world} fused_ms={t_fused:.4f} nccl_allgather_plus_attn_ms={t_nccl:.4f}", flush=True)
    heap.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
