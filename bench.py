#!/usr/bin/env python
"""Headline benchmark: tasks/hour + p50 task latency for AnythingV3 (SD-1.5) 512x512 / 20 steps / CFG, with ZeroScope
v2 XL, Kandinsky 2.1, the stock-library comparator and the multi-GPU fused paths in the same JSON line.

A "step" is one complete mining task: tokenise -> text encoder -> UNet steps at batch 2 (classifier-free guidance) ->
sampler -> VAE decode -> uint8 (-> PNG / MP4 + CIDv0 for the end-to-end number). `value` is AnythingV3 tasks/hour with N
GPUs running independent tasks (one process per GPU: weak scaling). The same line also carries

  eager          the same architecture on stock PyTorch kernels (cuDNN / cuBLAS / ATen / SDPA), timed in this run on this
                 box: plain NCHW at batch 1, and channels_last + cudnn.benchmark at the native arm's task batch;
                 `vs_baseline` = native / the better of the two (the reference itself cannot run: --impl reference)
  zeroscope_xl   the second headline model: 1024x576x24, 50 steps, native and eager
  kandinsky2     the reference's one enabled model, 768x768, 100 steps
  latency_mode   (N >= 2) ONE task over several GPUs — strong scaling: SD guidance-parallel pairs, ZeroScope
                 frame-parallel over N GPUs, Kandinsky patch-parallel over N GPUs — each with the CID compared against
                 the single-GPU run of the same task in this process
  roofline       tensor FLOPs per task (counted from the engine's call log) over the measured cuBLAS peak

  python bench.py --gpus 1 --steps 9 --warmup 3            # everything above (about five minutes)
  python bench.py --headline-only ...                      # AnythingV3 block only
  python bench.py --impl eager ...                         # stock-library PyTorch restatement as the main arm
  python bench.py --impl reference ...                     # the reference itself (unavailable: TS/Solidity)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

METRIC = "tasks/hour + p50 task latency (AnythingV3 512^2, 20 steps, CFG)"
ROOT = os.path.dirname(os.path.abspath(__file__))


def reference_arm() -> int:
    why = ("mainnet-pat/arbius has no Python package or GPU code (TypeScript miner + Solidity; models run in "
           "external Cog containers by docker digest): pip install of /root/reference fails "
           "('neither setup.py nor pyproject.toml'), node/docker/network are absent")
    if int(os.environ.get("RANK", "0")) == 0:               # one line per job, also when launched through torchrun
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [c.strip() for c in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for n, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks() -> dict:
    """The roofline denominators: driver-measured on this pool, else the profiling recipe's stated fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return {"bf16_tflops_sustained": float(d["bf16_tflops_sustained"]), "bf16_tflops": float(d["bf16_tflops"]),
                "hbm_gbs": float(d["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json)"}
    except (OSError, KeyError, ValueError):
        return {"bf16_tflops_sustained": 1400.0, "bf16_tflops": 1590.0, "hbm_gbs": 6650.0, "source": "fallback"}


NVLINK_GBS = 770.0           # measured peer copy per direction per GPU on this pool (B200_PROFILING.md)


class Run:
    """Process / rank context and the timing helpers every block uses."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = f"cuda:{self.local}"
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.dt = torch.float16 if args.dtype == "fp16" else torch.bfloat16

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def rmax(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_true(self, ok: bool) -> bool:
        return self.rmax(0.0 if ok else 1.0) == 0.0

    def timed(self, fn, n: int) -> float:
        """Device time of n calls of fn(i): barrier + synchronize on both sides, CUDA events on the launching stream,
        max over ranks. Returns ms per call."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        self.barrier()
        return self.rmax(e0.elapsed_time(e1)) / n

    def wall(self, fn) -> float:
        """End-to-end wall time of fn() in seconds (public API incl. H2D / D2H / file encode), max over ranks."""
        self.barrier()
        t0 = time.perf_counter()
        fn()
        self.torch.cuda.synchronize()
        s = self.rmax(time.perf_counter() - t0)
        self.barrier()
        return s

    def free(self):
        import gc
        gc.collect()
        self.torch.cuda.empty_cache()


def record_flops(fn) -> float:
    """TFLOP of tensor-core work issued by fn(), counted from the engine's call log."""
    from arbius_b200.utils.flops import total_flops
    g = sys.modules["arbius_b200.ops.gemm"]
    g.RECORD = []
    try:
        fn()
        return total_flops(g.RECORD) / 1e12
    finally:
        g.RECORD = None


# ------------------------------------------------------------------------------------------------ AnythingV3 (headline)
def sd_block(R: Run, impl: str, K: int, W: int, TB: int, model: str, size: int, dsteps: int) -> dict:
    torch = R.torch
    from arbius_b200 import ops
    from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline
    from arbius_b200.protocol import solution_cid
    cfg = SD15Config() if model == "anythingv3" else SD15Config.tiny()
    pipe = SD15Pipeline(cfg, backend=impl, device=R.dev, dtype=R.dt, seed=0)
    base = dict(prompt="arbius test cat, masterpiece, best quality", negative_prompt="lowres, bad anatomy",
                width=size, height=size, num_inference_steps=dsteps, guidance_scale=12.0, scheduler="DPMSolverMultistep")
    TB = TB if impl == "native" else 1
    counted = [0]

    def run_tasks(seeds):
        evs, i = [], 0
        while i < len(seeds):
            n = min(TB, len(seeds) - i)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            if n == 1:
                pipe.generate(seed=seeds[i], **base)
            else:
                pipe.generate_batch([dict(base, seed=sd) for sd in seeds[i:i + n]])
            b.record()
            counted[0] += getattr(pipe, "last_launches", 0)
            evs += [(a, b)] * n
            i += n
        return evs

    # ---- warm-up (graph capture happens here); the first pass also counts the task's tensor FLOPs ----
    task_tflop = None
    if impl == "native":
        tf = record_flops(lambda: pipe.generate(seed=999 + R.rank, **base))
        task_tflop = tf / 2.0                        # the first call runs the task twice: eager warm-up + capture pass
    for i in range(W):
        pipe.generate(seed=1000 + R.rank * 100 + i, **base)
    if TB > 1:
        for i in range(W):
            run_tasks([1500 + R.rank * 100 + i * TB + j for j in range(TB)])
            if K % TB > 1:
                run_tasks([1700 + R.rank * 100 + i * TB + j for j in range(K % TB)])
    R.barrier()
    n0 = ops.native.launch_count() if impl == "native" else 0
    sampler = ClockSampler(R.local).start()
    R.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    counted[0] = 0
    lat = run_tasks([2000 + R.rank * 1000 + i for i in range(K)])
    ev1.record()
    R.barrier()
    dev_ms = R.rmax(ev0.elapsed_time(ev1))
    lat_ms = sorted(a.elapsed_time(b) for a, b in lat)
    clocks = sampler.stop()
    launches = counted[0] if impl == "native" else 0
    eager_launches = (ops.native.launch_count() - n0) if impl == "native" else 0

    single = None
    if TB > 1:
        single_ms = R.timed(lambda i: pipe.generate(seed=2000 + R.rank * 1000 + i, **base), K)
        single = {"task_batch": 1, "ms_per_task": round(single_ms, 3), "p50_latency_ms": round(single_ms, 3),
                  "tasks_per_hour": round(R.world / (single_ms / 1e3) * 3600.0, 2)}

    cid = [None]
    e2e_inputs = [dict(base, seed=3000 + R.rank * 1000 + i) for i in range(K)]

    def e2e():
        if hasattr(pipe, "predict_many"):
            blobs = pipe.predict_many(e2e_inputs, task_batch=TB) if TB > 1 else pipe.predict_many(e2e_inputs)
        else:
            blobs = [pipe.predict(inp) for inp in e2e_inputs]
        for blob in blobs:
            cid[0] = solution_cid({"out-1.png": blob})
    e2e_s = R.wall(e2e)
    e2e_group = (K % TB or TB) if TB > 1 else 1
    tph = R.world * K / (dev_ms / 1e3) * 3600.0
    p50_replay = lat_ms[len(lat_ms) // 2]
    out = {
        "value": round(tph, 2), "ms_per_step": round(dev_ms / K, 3),
        # under batching one replay carries TB tasks: p50_replay_ms is the replay's device time; a task that arrives
        # uniformly within the previous replay waits half a replay more on average (p50_task_latency_ms)
        "p50_latency_ms": round(p50_replay, 3), "p50_replay_ms": round(p50_replay, 3),
        "p50_task_latency_ms": round(p50_replay * (1.5 if TB > 1 else 1.0), 3),
        "clocks": clocks,
        "e2e": {"value": round(R.world * K / e2e_s * 3600.0, 2), "unit": "tasks/hour", "ms_per_task": round(e2e_s / K * 1e3, 3),
                "h2d_bytes_per_step": pipe.h2d_bytes // e2e_group, "d2h_bytes_per_step": pipe.d2h_bytes // e2e_group,
                "includes": "per task: tokenise, pinned H2D ids, Philox noise, graph replay, D2H pixels, PNG encode, CIDv0"
                            " (predict_many: encode of task i overlaps the GPU work of task i+1)",
                "last_cid": cid[0]},
        "gpu_launches": int(launches or eager_launches), "task_batch": TB, "single_task": single, "task_tflop": task_tflop,
    }
    del pipe
    R.free()
    return out


def sd_eager_comparator(R: Run, TB: int, size: int, dsteps: int, K: int = 3) -> dict:
    """Stock-library arm timed in the same invocation on the same box and clocks: (a) plain NCHW at batch 1 — what the
    reference's containers run — and (b) the strongest stock configuration at an equal task batch: channels_last,
    cudnn.benchmark, UNet batch 2*TB."""
    torch = R.torch
    from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline
    from arbius_b200.protocol import solution_cid
    pipe = SD15Pipeline(SD15Config(), backend="eager", device=R.dev, dtype=R.dt, seed=0)
    base = dict(prompt="arbius test cat, masterpiece, best quality", negative_prompt="lowres, bad anatomy",
                width=size, height=size, num_inference_steps=dsteps, guidance_scale=12.0, scheduler="DPMSolverMultistep")
    for i in range(2):
        pipe.generate(seed=10 + i, **base)
    sampler = ClockSampler(R.local).start()
    nchw_ms = R.timed(lambda i: pipe.generate(seed=20 + i, **base), K)
    e2e_s = R.wall(lambda: [solution_cid({"out-1.png": pipe.predict(dict(base, seed=30 + i))}) for i in range(K)])
    torch.backends.cudnn.benchmark = True
    tasks = lambda s0: [dict(base, seed=s0 + j) for j in range(TB)]
    for i in range(3):
        pipe.generate_batch_eager(tasks(40 + 10 * i), channels_last=True)
    cl_ms = R.timed(lambda i: pipe.generate_batch_eager(tasks(100 + 10 * i), channels_last=True), K) / TB
    clocks = sampler.stop()
    torch.backends.cudnn.benchmark = False
    del pipe
    R.free()
    best = min(nchw_ms, cl_ms)
    return {"impl": "eager (stock PyTorch: cuDNN conv, cuBLAS, ATen GroupNorm, SDPA)", "timed_tasks": K,
            "nchw_batch1": {"ms_per_task": round(nchw_ms, 3), "tasks_per_hour": round(R.world * 3.6e6 / nchw_ms, 2)},
            "channels_last_cudnn_benchmark_equal_batch": {"task_batch": TB, "ms_per_task": round(cl_ms, 3),
                                                           "tasks_per_hour": round(R.world * 3.6e6 / cl_ms, 2)},
            "best_tasks_per_hour": round(R.world * 3.6e6 / best, 2),
            "e2e_tasks_per_hour": round(R.world * K / e2e_s * 3600.0, 2), "clocks": clocks}


# ------------------------------------------------------------------------------------------------ text-to-video
def video_cfg(model: str):
    from arbius_b200.models.video.pipeline import TextToVideoConfig
    return {"zeroscopev2xl": TextToVideoConfig.zeroscope_xl, "damo": TextToVideoConfig.damo,
            "video-tiny": TextToVideoConfig.tiny}[model]()


def video_base(vcfg, dsteps: int) -> dict:
    return dict(prompt="a panda eating bamboo on a rock", negative_prompt="noisy, washed out, ugly, distorted, broken",
                num_frames=vcfg.default_frames, num_inference_steps=dsteps, width=vcfg.default_size[0],
                height=vcfg.default_size[1], guidance_scale=vcfg.default_guidance)


def video_block(R: Run, model: str, impl: str, K: int, W: int, dsteps: int, with_e2e: bool = True, flops: bool = True) -> dict:
    from arbius_b200.models.video.pipeline import TextToVideoPipeline
    from arbius_b200.protocol import solution_cid
    vcfg = video_cfg(model)
    pipe = TextToVideoPipeline(vcfg, backend=impl, device=R.dev, dtype=R.dt, family=model)
    base = video_base(vcfg, dsteps)
    task_tflop = None
    if impl == "native" and flops:
        task_tflop = record_flops(lambda: pipe.generate(seed=900 + R.rank, **base)) / 2.0
    for i in range(W):
        pipe.generate(seed=1000 + R.rank * 100 + i, **base)
    sampler = ClockSampler(R.local).start()
    ms = R.timed(lambda i: pipe.generate(seed=2000 + R.rank * 1000 + i, **base), K)
    clocks = sampler.stop()
    out = {"model": f"{model} UNet3D {vcfg.default_size[0]}x{vcfg.default_size[1]}x{vcfg.default_frames}, {dsteps} steps",
           "impl": impl, "steps": K, "warmup": W, "ms_per_task": round(ms, 2), "p50_latency_ms": round(ms, 2),
           "tasks_per_hour": round(R.world * 3.6e6 / ms, 2), "clocks": clocks, "task_tflop": task_tflop,
           "gpu_launches": int(getattr(pipe, "last_launches", 0)) * K if impl == "native" else 0}
    if with_e2e:
        cid = [None]

        def e2e():
            cid[0] = solution_cid({"out-1.mp4": pipe.predict(dict(base, seed=3000 + R.rank, fps=vcfg.default_fps))})
        s = R.wall(e2e)
        out["e2e"] = {"value": round(R.world * 3600.0 / s, 2), "unit": "tasks/hour", "ms_per_task": round(s * 1e3, 1),
                      "h2d_bytes_per_step": pipe.h2d_bytes, "d2h_bytes_per_step": pipe.d2h_bytes, "last_cid": cid[0]}
    del pipe
    R.free()
    return out


# ------------------------------------------------------------------------------------------------ Kandinsky 2.1
def kandinsky_block(R: Run, model: str, impl: str, K: int, W: int) -> dict:
    from arbius_b200.models.kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
    from arbius_b200.protocol import solution_cid
    kcfg = Kandinsky2Config() if model == "kandinsky2" else Kandinsky2Config.tiny()
    ksz = 768 if model == "kandinsky2" else 128
    pipe = Kandinsky2Pipeline(kcfg, backend=impl, device=R.dev, dtype=R.dt)
    task_tflop = None
    if impl == "native":
        task_tflop = record_flops(lambda: pipe.generate("arbius test cat", ksz, ksz, seed=900 + R.rank)) / 2.0
    for i in range(W):
        pipe.generate("arbius test cat", ksz, ksz, seed=1000 + i)
    sampler = ClockSampler(R.local).start()
    ms = R.timed(lambda i: pipe.generate("arbius test cat", ksz, ksz, seed=2000 + R.rank * 1000 + i), K)
    clocks = sampler.stop()
    cid = [None]

    def e2e():
        for i in range(K):
            cid[0] = solution_cid({"out-1.png": pipe.predict({"prompt": "arbius test cat", "width": ksz, "height": ksz,
                                                               "seed": 3000 + R.rank * 1000 + i})})
    s = R.wall(e2e)
    out = {"model": f"{model} (XLM-R + diffusion prior + UNet 1.2B + MoVQ) {ksz}x{ksz}, {kcfg.steps} steps", "impl": impl,
           "steps": K, "warmup": W, "ms_per_task": round(ms, 2), "p50_latency_ms": round(ms, 2),
           "tasks_per_hour": round(R.world * 3.6e6 / ms, 2), "clocks": clocks, "task_tflop": task_tflop,
           "gpu_launches": int(getattr(pipe, "last_launches", 0)) * K if impl == "native" else 0,
           "e2e": {"value": round(R.world * K / s * 3600.0, 2), "unit": "tasks/hour", "ms_per_task": round(s / K * 1e3, 2),
                   "h2d_bytes_per_step": pipe.h2d_bytes, "d2h_bytes_per_step": pipe.d2h_bytes, "last_cid": cid[0]}}
    del pipe
    R.free()
    return out


# ------------------------------------------------------------------------------------------------ ONE task over N GPUs
def latency_mode(R: Run, K: int, W: int, do_sd: bool, do_video, do_kandinsky, dsteps_sd: int, size: int) -> dict:
    """Strong scaling: the same single task spread over GPUs, every path compared byte for byte (CID of the encoded
    file) with the single-GPU pipeline of THIS process on the same seed. Device-timed, max over ranks."""
    torch = R.torch
    from arbius_b200.parallel.groups import subset_group
    from arbius_b200.parallel.symm import PeerHeap
    from arbius_b200.protocol import solution_cid
    out = {}
    world = R.world
    peaks = measured_peaks()
    if do_sd and world % 2 == 0:
        from arbius_b200.codecs.png import encode_png
        from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline
        heap = PeerHeap(64 << 20, group=subset_group(2))
        base = dict(prompt="arbius test cat, masterpiece, best quality", negative_prompt="lowres, bad anatomy",
                    width=size, height=size, num_inference_steps=dsteps_sd, guidance_scale=12.0, scheduler="DPMSolverMultistep")
        one = SD15Pipeline(SD15Config(), backend="native", device=R.dev, dtype=R.dt, seed=0)
        par = SD15Pipeline(SD15Config(), backend="native", device=R.dev, dtype=R.dt, seed=0, cfg_parallel=heap)
        pair = R.rank // 2
        ref_cid = solution_cid({"out-1.png": encode_png(one.generate(seed=77, **base))})
        for i in range(W):
            par.generate(seed=5000 + pair * 100 + i, **base)
        same = solution_cid({"out-1.png": encode_png(par.generate(seed=77, **base))}) == ref_cid
        ms1 = R.timed(lambda i: one.generate(seed=6000 + i, **base), K)
        ms = R.timed(lambda i: par.generate(seed=6000 + pair * 1000 + i, **base), K)
        out["anythingv3_guidance_parallel"] = {
            "parallelism": f"guidance-parallel pairs x task-dp{world // 2}", "gpus_per_task": 2,
            "ms_per_task": round(ms, 3), "single_gpu_ms_per_task": round(ms1, 3), "speedup": round(ms1 / ms, 3),
            "tasks_per_hour": round((world // 2) * 3.6e6 / ms, 2), "cid_equal_to_p1": R.all_true(same),
            "nvlink_bytes_per_task": 2 * dsteps_sd * (size // 8) ** 2 * 4 * 2}
        heap.close()
        del one, par
        R.free()
    if do_video:
        from arbius_b200.codecs.mp4 import encode_mp4
        from arbius_b200.models.video.pipeline import TextToVideoPipeline, frame_split_ok
        model, dsteps = do_video
        vcfg = video_cfg(model)
        base = video_base(vcfg, dsteps)
        h8, w8 = vcfg.default_size[1] // 8, vcfg.default_size[0] // 8
        if frame_split_ok(vcfg.default_frames, h8, w8, world):
            one = TextToVideoPipeline(vcfg, backend="native", device=R.dev, dtype=R.dt, family=model)
            ref = one.generate(seed=77, **base)
            ref_cid = solution_cid({"out-1.mp4": encode_mp4(ref, vcfg.default_fps)})
            ms1 = R.timed(lambda i: one.generate(seed=6000 + i, **base), 1)
            tflop = record_flops(lambda: None)
            del one, ref
            R.free()
            heap = PeerHeap((3 << 30) if model != "video-tiny" else (256 << 20))
            par = TextToVideoPipeline(vcfg, backend="native", device=R.dev, dtype=R.dt, family=model, frame_parallel=heap)
            for i in range(max(1, W - 2)):
                par.generate(seed=5000 + i, **base)
            got = par.generate(seed=77, **base)
            same = True if got is None else solution_cid({"out-1.mp4": encode_mp4(got, vcfg.default_fps)}) == ref_cid
            ms = R.timed(lambda i: par.generate(seed=6000 + i, **base), max(1, min(K, 2)))
            out[f"{model}_frame_parallel"] = {
                "parallelism": f"frame-parallel sp{world} (temporal GroupNorm / conv / attention cross ranks through peer memory)",
                "gpus_per_task": world, "denoise_steps": dsteps, "ms_per_task": round(ms, 2),
                "single_gpu_ms_per_task": round(ms1, 2), "speedup": round(ms1 / ms, 3),
                "tasks_per_hour": round(3.6e6 / ms, 2), "cid_equal_to_p1": R.all_true(same)}
            heap.close()
            del par
            R.free()
    if do_kandinsky and world in (2, 4, 8):
        from arbius_b200.codecs.png import encode_png
        from arbius_b200.models.kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
        model = do_kandinsky
        kcfg = Kandinsky2Config() if model == "kandinsky2" else Kandinsky2Config.tiny()
        ksz = 768 if model == "kandinsky2" else 128
        one = Kandinsky2Pipeline(kcfg, backend="native", device=R.dev, dtype=R.dt)
        ref_cid = solution_cid({"out-1.png": encode_png(one.generate("arbius test cat", ksz, ksz, seed=77))})
        tflop = record_flops(lambda: one._run_native(one._graphs[(ksz, ksz)]))
        ms1 = R.timed(lambda i: one.generate("arbius test cat", ksz, ksz, seed=6000 + i), K)
        del one
        R.free()
        heap = PeerHeap((6 << 30) if model == "kandinsky2" else (512 << 20))
        par = Kandinsky2Pipeline(kcfg, backend="native", device=R.dev, dtype=R.dt, patch_parallel=heap)
        for i in range(W):
            par.generate("arbius test cat", ksz, ksz, seed=5000 + i)
        same = solution_cid({"out-1.png": encode_png(par.generate("arbius test cat", ksz, ksz, seed=77))}) == ref_cid
        ms = R.timed(lambda i: par.generate("arbius test cat", ksz, ksz, seed=6000 + i), K)
        # roofline of the fused path: the slower of the task's tensor FLOPs / P at the measured GEMM peak and the bytes
        # one GPU pulls over NVLink (the other bands' K/V of every attention layer, every step) at the measured link rate
        lat = ksz // 8
        kv_bytes = 0
        for lvl_hw, ch, n_att in (((lat // 2) ** 2, 768, 7), ((lat // 4) ** 2, 1152, 7)) if model == "kandinsky2" else ():
            kv_bytes += n_att * 2 * 2 * lvl_hw * ch * 2 * (world - 1) // world
        kv_bytes *= kcfg.steps
        t_flop_ms = tflop / world / peaks["bf16_tflops_sustained"] * 1e3
        t_link_ms = kv_bytes / (NVLINK_GBS * 1e9) * 1e3
        bound = max(t_flop_ms, t_link_ms)
        out[f"{model}_patch_parallel"] = {
            "parallelism": f"patch-parallel pp{world} (row bands; GroupNorm moments, conv halo rows and attention K/V cross "
                           "ranks through peer memory with in-kernel flag waits)",
            "gpus_per_task": world, "ms_per_task": round(ms, 2), "single_gpu_ms_per_task": round(ms1, 2),
            "speedup": round(ms1 / ms, 3), "tasks_per_hour": round(3.6e6 / ms, 2), "cid_equal_to_p1": R.all_true(same),
            "gpu_launches_per_task": int(par.last_launches),
            "roofline": {"task_tflop": round(tflop, 2), "nvlink_bytes_per_gpu_per_task": int(kv_bytes),
                         "compute_bound_ms": round(t_flop_ms, 2), "nvlink_bound_ms": round(t_link_ms, 2),
                         "achieved_fraction_of_roofline": round(bound / ms, 4) if ms > 0 else None,
                         "denominators": f"bf16 sustained {peaks['bf16_tflops_sustained']} TFLOP/s ({peaks['source']}), "
                                         f"NVLink {NVLINK_GBS} GB/s per direction (measured peer copy)"}}
        heap.close()
        del par
        R.free()
    return out


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=9, help="timed tasks per GPU (headline block)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "eager", "reference"])
    ap.add_argument("--model", default="anythingv3", choices=["anythingv3", "tiny", "zeroscopev2xl", "damo", "video-tiny",
                                                              "kandinsky2", "kandinsky-tiny"])
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--denoise-steps", type=int, default=20)
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--headline-only", action="store_true", help="skip the comparator, ZeroScope, Kandinsky and latency blocks")
    ap.add_argument("--latency-mode", action="store_true", help="(kept for compatibility: latency mode is on by default at N >= 2)")
    ap.add_argument("--no-latency-mode", action="store_true")
    ap.add_argument("--task-batch", type=int, default=4,
                    help="independent tasks per graph replay (anythingv3 native only); bytes per task do not change")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm()
    R = Run(args)
    W, K = max(args.warmup, 3), args.steps
    peaks = measured_peaks()
    full = args.model == "anythingv3" and args.impl == "native" and not args.headline_only

    if args.model in ("anythingv3", "tiny"):
        sd = sd_block(R, args.impl, K, W, args.task_batch, args.model, args.size, args.denoise_steps)
        TB = sd["task_batch"]
        model_desc = "AnythingV3 (SD-1.5 UNet 860M + CLIP ViT-L/14 + AutoencoderKL)" if args.model == "anythingv3" else "tiny"
        out = {"metric": METRIC, "value": sd["value"], "unit": "tasks/hour", "impl": args.impl, "n_gpus": R.world, "steps": K,
               "warmup": W, "ms_per_step": sd["ms_per_step"], "p50_latency_ms": sd["p50_latency_ms"],
               "p50_replay_ms": sd["p50_replay_ms"], "p50_task_latency_ms": sd["p50_task_latency_ms"],
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": args.dtype,
               "data": "synthetic prompts, random-init weights of the SD-1.5 architecture",
               "config": {"model": model_desc, "global_batch": R.world * TB, "seq_len": 77, "image": f"{args.size}x{args.size}",
                          "denoise_steps": args.denoise_steps, "scheduler": "DPMSolverMultistep", "cfg_batch": 2,
                          "task_batch": TB, "parallelism": f"task-dp{R.world}",
                          "l2": "inputs larger than L2: every task streams 2.06 GB of weights through a 126 MB L2"},
               "clocks": sd["clocks"], "e2e": sd["e2e"], "gpu_launches": sd["gpu_launches"]}
        if sd["single_task"] is not None:
            out["single_task"] = sd["single_task"]
            out["config"]["batching"] = (f"{TB} independent tasks per CUDA-graph replay (UNet batch {2 * TB}); each task's bytes "
                                         "equal the ones it gets alone (tests/test_sd15.py); p50_latency_ms is the replay's "
                                         "device time, p50_task_latency_ms adds the mean wait of a task arriving mid-replay")
        roof = {"denominator": f"bf16 sustained {peaks['bf16_tflops_sustained']} TFLOP/s ({peaks['source']})"}
        if sd["task_tflop"]:
            per_gpu_tps = sd["value"] / 3600.0 / R.world
            roof["anythingv3"] = {"task_tflop": round(sd["task_tflop"], 2),
                                  "achieved_tflops_per_gpu": round(sd["task_tflop"] * per_gpu_tps, 1),
                                  "fraction_of_measured_peak": round(sd["task_tflop"] * per_gpu_tps / peaks["bf16_tflops_sustained"], 4)}
        out["roofline"] = roof
        if full:
            eg = sd_eager_comparator(R, TB, args.size, args.denoise_steps)
            out["eager"] = {"anythingv3": eg}
            out["vs_baseline"] = round(out["value"] / eg["best_tasks_per_hour"], 3)
            out["vs_baseline_note"] = ("native tasks/hour / the better stock-PyTorch arm timed in this run (no published reference "
                                       "number exists and the reference cannot run offline: BASELINE.md)")
            out["e2e"]["vs_baseline"] = round(out["e2e"]["value"] / eg["e2e_tasks_per_hour"], 3)
            if out.get("single_task"):
                out["single_task"]["vs_eager_nchw_batch1"] = round(eg["nchw_batch1"]["ms_per_task"] / out["single_task"]["ms_per_task"], 3)
            z = video_block(R, "zeroscopev2xl", "native", 2, 1, 50)
            ze = video_block(R, "zeroscopev2xl", "eager", 1, 1, 50, with_e2e=False)
            z["eager"] = {"ms_per_task": ze["ms_per_task"], "tasks_per_hour": ze["tasks_per_hour"], "clocks": ze["clocks"]}
            z["vs_baseline"] = round(z["tasks_per_hour"] / ze["tasks_per_hour"], 3)
            out["zeroscope_xl"] = z
            if z.get("task_tflop"):
                tps = z["tasks_per_hour"] / 3600.0 / R.world
                roof["zeroscope_xl"] = {"task_tflop": round(z["task_tflop"], 1), "achieved_tflops_per_gpu": round(z["task_tflop"] * tps, 1),
                                        "fraction_of_measured_peak": round(z["task_tflop"] * tps / peaks["bf16_tflops_sustained"], 4)}
            kd = kandinsky_block(R, "kandinsky2", "native", 3, 3)
            out["kandinsky2"] = kd
            if kd.get("task_tflop"):
                tps = kd["tasks_per_hour"] / 3600.0 / R.world
                roof["kandinsky2"] = {"task_tflop": round(kd["task_tflop"], 1), "achieved_tflops_per_gpu": round(kd["task_tflop"] * tps, 1),
                                      "fraction_of_measured_peak": round(kd["task_tflop"] * tps / peaks["bf16_tflops_sustained"], 4)}
            if R.world >= 2 and not args.no_latency_mode:
                out["latency_mode"] = latency_mode(R, 3, W, True, ("zeroscopev2xl", 50), "kandinsky2", args.denoise_steps, args.size)
        elif args.model == "tiny" and R.world >= 2 and args.impl == "native" and not args.no_latency_mode and not args.headline_only:
            out["latency_mode"] = latency_mode(R, 2, W, False, ("video-tiny", 3), "kandinsky-tiny", args.denoise_steps, args.size)
    elif args.model in ("zeroscopev2xl", "damo", "video-tiny"):
        dsteps = args.denoise_steps if args.denoise_steps != 20 else 50          # template default: 50 steps
        v = video_block(R, args.model, args.impl, K, min(W, 3), dsteps)
        out = {"metric": f"tasks/hour + p50 task latency ({v['model']})", "value": v["tasks_per_hour"], "unit": "tasks/hour",
               "impl": args.impl, "n_gpus": R.world, "steps": K, "warmup": min(W, 3), "ms_per_step": v["ms_per_task"],
               "p50_latency_ms": v["p50_latency_ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": args.dtype, "data": "synthetic prompts, random-init weights of the text-to-video UNet3D architecture",
               "config": {"model": v["model"], "global_batch": R.world, "seq_len": 77, "cfg_batch": 2, "parallelism": f"task-dp{R.world}",
                          "l2": "inputs larger than L2 (multi-GB activations and weights per step)"},
               "clocks": v["clocks"], "e2e": v.get("e2e"), "gpu_launches": v["gpu_launches"]}
        if R.world >= 2 and args.impl == "native" and not args.no_latency_mode:
            out["latency_mode"] = latency_mode(R, K, W, False, (args.model, dsteps), None, 20, 512)
    else:
        kd = kandinsky_block(R, args.model, args.impl, K, W)
        out = {"metric": f"tasks/hour + p50 task latency ({kd['model']})", "value": kd["tasks_per_hour"], "unit": "tasks/hour",
               "impl": args.impl, "n_gpus": R.world, "steps": K, "warmup": W, "ms_per_step": kd["ms_per_task"],
               "p50_latency_ms": kd["p50_latency_ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": args.dtype, "data": "synthetic prompts, random-init weights of the Kandinsky 2.1 architecture",
               "config": {"model": kd["model"], "global_batch": R.world, "seq_len": 77, "cfg_batch": 2, "parallelism": f"task-dp{R.world}",
                          "l2": "inputs larger than L2: every step streams 2.4 GB of weights"},
               "clocks": kd["clocks"], "e2e": kd["e2e"], "gpu_launches": kd["gpu_launches"]}
        if R.world >= 2 and args.impl == "native" and not args.no_latency_mode:
            out["latency_mode"] = latency_mode(R, K, W, False, None, args.model, 20, 512)
    if R.rank == 0:
        print(json.dumps(out))
    if R.world > 1:
        R.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
