"""torchrun worker (2 ranks): guidance-parallel SD-1.5 must reproduce the single-GPU bytes, also through graph replay."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline  # noqa: E402
from arbius_b200.parallel import init_distributed  # noqa: E402
from arbius_b200.parallel.symm import PeerHeap  # noqa: E402


def main():
    world = init_distributed()
    assert world == 2, "guidance parallelism uses exactly two ranks"
    rank = dist.get_rank()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    heap = PeerHeap(64 << 20)
    ok = True
    ref = SD15Pipeline(SD15Config.tiny(), backend="native", device=dev, dtype=torch.float16)
    par = SD15Pipeline(SD15Config.tiny(), backend="native", device=dev, dtype=torch.float16, cfg_parallel=heap)
    for sched in ("DPMSolverMultistep", "K_EULER_ANCESTRAL"):
        kw = dict(prompt="arbius test cat", negative_prompt="blurry", width=64, height=64, num_inference_steps=4,
                  guidance_scale=7.5, scheduler=sched)
        a = ref.generate(seed=1337, **kw)
        b1 = par.generate(seed=1337, **kw)            # warm-up + capture + first replay
        b2 = par.generate(seed=1337, **kw)            # replay: the barrier epochs must keep advancing
        c = par.generate(seed=7, **kw)
        same = bool((a == b1).all() and (a == b2).all() and (c != a).any())
        print(f"rank {rank} {sched} bit-identical={same}", flush=True)
        ok = ok and same
    if os.environ.get("AB_CFGP_FULL", "1") == "1":
        full_ref = SD15Pipeline(SD15Config(), backend="native", device=dev, dtype=torch.float16)
        full_par = SD15Pipeline(SD15Config(), backend="native", device=dev, dtype=torch.float16, cfg_parallel=heap)
        kw = dict(prompt="arbius test cat", negative_prompt="lowres", width=512, height=512, num_inference_steps=20,
                  guidance_scale=12.0, scheduler="DPMSolverMultistep")
        a = full_ref.generate(seed=3, **kw)
        b = full_par.generate(seed=3, **kw)
        same = bool((a == b).all())
        print(f"rank {rank} AnythingV3 512x512 bit-identical={same}", flush=True)
        ok = ok and same

        def timeit(p, n=4):
            for _ in range(2):
                p.generate(seed=5, **kw)
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            for i in range(n):
                p.generate(seed=10 + i, **kw)
            torch.cuda.synchronize()
            t = torch.tensor([(time.perf_counter() - t0) / n * 1e3], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()
        t_par = timeit(full_par)
        t_ref = timeit(full_ref)
        if rank == 0:
            print(f"RESULT cfg_parallel world=2 latency_ms={t_par:.2f} single_gpu_latency_ms={t_ref:.2f}", flush=True)
    # ---------------- Kandinsky 2.1 ----------------
    from arbius_b200.models.kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
    kref = Kandinsky2Pipeline(Kandinsky2Config.tiny(), backend="native", device=dev, dtype=torch.float16)
    kpar = Kandinsky2Pipeline(Kandinsky2Config.tiny(), backend="native", device=dev, dtype=torch.float16, cfg_parallel=heap)
    a = kref.generate("arbius test cat", 32, 32, seed=5)
    b1 = kpar.generate("arbius test cat", 32, 32, seed=5)
    b2 = kpar.generate("arbius test cat", 32, 32, seed=5)
    same = bool((a == b1).all() and (a == b2).all())
    print(f"rank {rank} kandinsky-tiny bit-identical={same}", flush=True)
    ok = ok and same
    if os.environ.get("AB_CFGP_FULL", "1") == "1":
        del kref, kpar
        torch.cuda.empty_cache()
        kref = Kandinsky2Pipeline(Kandinsky2Config(), backend="native", device=dev, dtype=torch.float16)
        kpar = Kandinsky2Pipeline(Kandinsky2Config(), backend="native", device=dev, dtype=torch.float16, cfg_parallel=heap)
        a = kref.generate("arbius test cat", 768, 768, seed=5)
        b = kpar.generate("arbius test cat", 768, 768, seed=5)
        same = bool((a == b).all())
        print(f"rank {rank} Kandinsky 2.1 768x768 bit-identical={same}", flush=True)
        ok = ok and same

        def ktime(p, n=3):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            for i in range(n):
                p.generate("arbius test cat", 768, 768, seed=30 + i)
            torch.cuda.synchronize()
            t = torch.tensor([(time.perf_counter() - t0) / n * 1e3], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()
        t_par, t_ref = ktime(kpar), ktime(kref)
        if rank == 0:
            print(f"RESULT kandinsky2 cfg_parallel world=2 latency_ms={t_par:.1f} single_gpu_latency_ms={t_ref:.1f}", flush=True)
        del kref, kpar
        torch.cuda.empty_cache()
    # ---------------- text-to-video UNet3D ----------------
    from arbius_b200.models.video.pipeline import TextToVideoConfig, TextToVideoPipeline
    vref = TextToVideoPipeline(TextToVideoConfig.tiny(), backend="native", device=dev, dtype=torch.float16, family="video-tiny")
    vpar = TextToVideoPipeline(TextToVideoConfig.tiny(), backend="native", device=dev, dtype=torch.float16, family="video-tiny",
                               cfg_parallel=heap)
    vkw = dict(prompt="a panda eating bamboo", negative_prompt="noisy", num_inference_steps=3)
    a = vref.generate(seed=11, **vkw)
    b1 = vpar.generate(seed=11, **vkw)
    b2 = vpar.generate(seed=11, **vkw)
    same = bool((a == b1).all() and (a == b2).all())
    print(f"rank {rank} video-tiny bit-identical={same}", flush=True)
    ok = ok and same
    if os.environ.get("AB_CFGP_FULL", "1") == "1":
        del full_ref, full_par
        torch.cuda.empty_cache()
        zcfg = TextToVideoConfig.zeroscope_xl()
        zref = TextToVideoPipeline(zcfg, backend="native", device=dev, dtype=torch.float16)
        zpar = TextToVideoPipeline(zcfg, backend="native", device=dev, dtype=torch.float16, cfg_parallel=heap)
        zkw = dict(prompt="a panda eating bamboo on a rock", negative_prompt="noisy, washed out", num_inference_steps=4)
        a = zref.generate(seed=3, **zkw)
        b = zpar.generate(seed=3, **zkw)
        same = bool((a == b).all())
        print(f"rank {rank} ZeroScope XL 1024x576x24 (4 steps) bit-identical={same}", flush=True)
        ok = ok and same

        def vtime(p):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            p.generate(seed=21, **zkw)
            torch.cuda.synchronize()
            t = torch.tensor([(time.perf_counter() - t0) * 1e3], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()
        t_par, t_ref = vtime(zpar), vtime(zref)
        if rank == 0:
            print(f"RESULT zeroscope cfg_parallel world=2 4-step latency_ms={t_par:.1f} single_gpu_latency_ms={t_ref:.1f}", flush=True)
    heap.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
