"""Per-kernel time of Kandinsky 2.1 on one GPU vs patch-parallel over the ranks of a torchrun job (CUPTI via
torch.profiler around one CUDA-graph replay; rank 0 prints both tables). Kernel times of the patch run include the
in-kernel flag waits, so the table shows where the hand-offs cost.

  python -m torch.distributed.run --nproc-per-node 2 scripts/profile_patch.py --steps 4
"""
import argparse
import os
import re
import sys
from collections import defaultdict

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def table(prof, title, top=28):
    agg = defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            full = ev.name
            key = re.sub(r"<.*", "", full)
            m = re.search(r"gemm_tc_kernel<(\d+)", full)
            if m:
                key = f"gemm_tc_kernel<BN={m.group(1)}>"
            agg[key][0] += ev.device_time
            agg[key][1] += 1
    total = sum(v[0] for v in agg.values())
    rows = sorted(([k, v[0] / 1e3, v[1]] for k, v in agg.items()), key=lambda r: -r[1])
    print(f"== {title}: total kernel time {total / 1e3:.2f} ms over {sum(r[2] for r in rows)} launches", flush=True)
    for k, ms, n in rows[:top]:
        print(f"{ms:9.3f} ms {100 * ms * 1e3 / total:5.1f}% {n:6d}x  avg {1e3 * ms / n:8.2f} us  {k}", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--size", type=int, default=768)
    a = ap.parse_args()
    from arbius_b200.models.kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
    from arbius_b200.parallel import init_distributed
    from arbius_b200.parallel.symm import PeerHeap
    from torch.profiler import ProfilerActivity, profile
    world = init_distributed()
    rank = dist.get_rank()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    cfg = Kandinsky2Config()
    cfg.steps = a.steps
    one = Kandinsky2Pipeline(cfg, backend="native", device=dev, dtype=torch.float16)
    for s in (1, 2):
        one.generate("arbius test cat", a.size, a.size, seed=s)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    one.generate("arbius test cat", a.size, a.size, seed=3)
    e1.record()
    torch.cuda.synchronize()
    t_one = e0.elapsed_time(e1)
    with profile(activities=[ProfilerActivity.CUDA]) as p1:
        one.generate("arbius test cat", a.size, a.size, seed=4)
        torch.cuda.synchronize()
    if rank == 0:
        table(p1, f"one GPU, {a.steps} UNet steps, {t_one:.1f} ms per task")
        p1.export_chrome_trace("gpurun_out/trace_kandinsky_p1.json")
    del one
    torch.cuda.empty_cache()
    heap = PeerHeap(6 << 30)
    par = Kandinsky2Pipeline(cfg, backend="native", device=dev, dtype=torch.float16, patch_parallel=heap)
    for s in (1, 2):
        par.generate("arbius test cat", a.size, a.size, seed=s)
    torch.cuda.synchronize()
    dist.barrier()
    e0.record()
    par.generate("arbius test cat", a.size, a.size, seed=3)
    e1.record()
    torch.cuda.synchronize()
    t_par = e0.elapsed_time(e1)
    dist.barrier()
    with profile(activities=[ProfilerActivity.CUDA]) as p2:
        par.generate("arbius test cat", a.size, a.size, seed=4)
        torch.cuda.synchronize()
    if rank == 0:
        table(p2, f"patch-parallel P={world}, {a.steps} UNet steps, {t_par:.1f} ms per task")
    p2.export_chrome_trace(f"gpurun_out/trace_kandinsky_patch_p{world}_rank{rank}.json")
    dist.barrier()
    heap.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
