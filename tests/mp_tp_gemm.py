"""torchrun worker: fused all-gather->GEMM and GEMM->reduce-scatter must equal the single-GPU kernels bit for bit."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arbius_b200 import ops  # noqa: E402
from arbius_b200.parallel import init_distributed  # noqa: E402
from arbius_b200.parallel.symm import PeerHeap  # noqa: E402
from arbius_b200.parallel.tp_gemm import AgGemm, GemmRs  # noqa: E402


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


def main():
    world = init_distributed()
    rank = dist.get_rank()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    heap = PeerHeap(512 << 20)
    dt = torch.float16
    ok = True
    # ---------------- all-gather -> GEMM (column parallel) ----------------
    for (m, n, k) in [(8192, 960, 320), (2048, 5120, 640), (1024, 1280, 1280)]:
        torch.manual_seed(7)
        x = torch.randn(m, k, device="cuda", dtype=dt)
        w = torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5
        b = torch.randn(n, device="cuda", dtype=dt)
        nl = n // world
        wl, bl = w[rank * nl:(rank + 1) * nl].contiguous(), b[rank * nl:(rank + 1) * nl].contiguous()
        ref = ops.gemm(x, wl, bl, act="silu", splitk=1, block_n=64)
        heap.reset()
        ag = AgGemm(heap, m // world, k, dt)
        ag.x_slot.copy_(x[rank * (m // world):(rank + 1) * (m // world)])
        out = ag(wl, bl, act="silu", block_n=64)
        torch.cuda.synchronize()
        same = torch.equal(out, ref)
        print(f"rank {rank} ag_gemm {(m, n, k)} bit-identical={same}", flush=True)
        ok = ok and same
    # ---------------- GEMM -> reduce-scatter (row parallel) ----------------
    for (m, n, k) in [(8192, 320, 1536), (2048, 640, 2560), (1024, 1280, 5120)]:     # K % 512 == 0
        torch.manual_seed(11)
        x = torch.randn(m, k, device="cuda", dtype=dt)
        w = torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5
        b = torch.randn(n, device="cuda", dtype=dt)
        res = torch.randn(m, n, device="cuda", dtype=dt)
        ref = ops.gemm(x, w, b, residual=res, splitk=8, block_n=64)       # canonical 8-shard order on one GPU
        kl, ml = k // world, m // world
        xl = x[:, rank * kl:(rank + 1) * kl].contiguous()
        wl = w[:, rank * kl:(rank + 1) * kl].contiguous()
        heap.reset()
        rs = GemmRs(heap, m, n)
        out = rs(xl, wl, b, residual=res[rank * ml:(rank + 1) * ml].contiguous(), block_n=64)
        torch.cuda.synchronize()
        same = torch.equal(out, ref[rank * ml:(rank + 1) * ml])
        err = (out.float() - ref[rank * ml:(rank + 1) * ml].float()).abs().max().item()
        print(f"rank {rank} gemm_rs {(m, n, k)} bit-identical={same} maxerr={err:.3e}", flush=True)
        ok = ok and same
    # ---------------- timing vs NCCL ----------------
    m, n, k = 8192, 5120, 1280
    x = torch.randn(m // world, k, device="cuda", dtype=dt)
    wl = torch.randn(n // world, k, device="cuda", dtype=dt)
    heap.reset()
    ag = AgGemm(heap, m // world, k, dt)
    ag.x_slot.copy_(x)

    def nccl_ag():
        full = torch.empty(m, k, device="cuda", dtype=dt)
        dist.all_gather_into_tensor(full, x)
        return ops.gemm(full, wl, splitk=1)
    t_f = timeit(lambda: ag(wl))
    t_n = timeit(nccl_ag)
    k = 5120                                              # FF-out of the 1280-wide level (K % 512 == 0)
    xk = torch.randn(m, k // world, device="cuda", dtype=dt)
    wk = torch.randn(1280, k // world, device="cuda", dtype=dt)

    def t_barrier():
        heap.barrier()
    t_b = timeit(t_barrier, iters=50)
    if rank == 0:
        print(f"INFO peer barrier (launch + NVLink round trip) {t_b * 1e3:.1f} us", flush=True)
    rs = GemmRs(heap, m, 1280)

    def nccl_rs():
        part = ops.gemm(xk, wk, splitk=1)
        out = torch.empty(m // world, 1280, device="cuda", dtype=dt)
        dist.reduce_scatter_tensor(out, part)
        return out
    t_rf = timeit(lambda: rs(xk, wk))
    t_rn = timeit(nccl_rs)
    if rank == 0:
        print(f"RESULT world={world} ag_gemm fused_ms={t_f:.4f} nccl_ms={t_n:.4f} | gemm_rs fused_ms={t_rf:.4f} nccl_ms={t_rn:.4f}", flush=True)
    heap.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
