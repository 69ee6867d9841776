"""Per-kernel time breakdown of one task (CUPTI via torch.profiler; shares, not absolutes).

  python scripts/profile_task.py                                   # AnythingV3 512x512, 20 steps
  python scripts/profile_task.py --model zeroscopev2xl --steps 2   # two denoise steps + VAE of the video task
"""
import argparse
import json
import re
import sys
from collections import defaultdict

import torch

sys.path.insert(0, ".")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="native")
    ap.add_argument("--model", default="anythingv3", choices=["anythingv3", "zeroscopev2xl", "damo", "kandinsky2"])
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--task-batch", type=int, default=1, help="anythingv3: independent tasks per forward")
    ap.add_argument("--out", default="gpurun_out/profile_task.json")
    a = ap.parse_args()
    if a.model == "anythingv3":
        from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline
        pipe = SD15Pipeline(SD15Config(), backend=a.impl, device="cuda", dtype=torch.float16, use_graph=False)
        kw = dict(prompt="arbius test cat", negative_prompt="", width=a.size, height=a.size,
                  num_inference_steps=a.steps, guidance_scale=12.0, scheduler="DPMSolverMultistep")
    elif a.model in ("zeroscopev2xl", "damo"):
        from arbius_b200.models.video.pipeline import TextToVideoConfig, TextToVideoPipeline
        cfg = TextToVideoConfig.zeroscope_xl() if a.model == "zeroscopev2xl" else TextToVideoConfig.damo()
        pipe = TextToVideoPipeline(cfg, backend=a.impl, device="cuda", dtype=torch.float16, family=a.model, use_graph=False)
        kw = dict(prompt="a panda eating bamboo", negative_prompt="noisy", num_inference_steps=a.steps)
    else:
        from arbius_b200.models.kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
        cfg = Kandinsky2Config()
        cfg.steps = a.steps
        pipe = Kandinsky2Pipeline(cfg, backend=a.impl, device="cuda", dtype=torch.float16, use_graph=False)
        kw = dict(prompt="arbius test cat", width=768, height=768)
    def run(seed):
        if a.task_batch > 1 and a.model == "anythingv3":
            pipe.generate_batch([dict(kw, seed=seed + i) for i in range(a.task_batch)])
        else:
            pipe.generate(seed=seed, **kw)
    run(1)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        run(10)
        torch.cuda.synchronize()
    agg = defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            full = ev.name
            key = re.sub(r"<.*", "", full)
            m = re.search(r"gemm_tc_kernel<(\d+)", full)
            if m:
                key = f"gemm_tc_kernel<BN={m.group(1)}>"
            m = re.search(r"flash_attn_kernel<(\d+)", full)
            if m:
                key = f"flash_attn_kernel<NCH={m.group(1)}>"
            agg[key][0] += ev.device_time
            agg[key][1] += 1
    total = sum(v[0] for v in agg.values())
    rows = sorted(([k, v[0] / 1e3, v[1]] for k, v in agg.items()), key=lambda r: -r[1])
    print(f"{a.model}: total kernel time {total / 1e3:.2f} ms over {sum(r[2] for r in rows)} launches")
    for k, ms, n in rows[:25]:
        print(f"{ms:9.3f} ms {100 * ms * 1e3 / total:5.1f}% {n:6d}x  avg {1e3 * ms / n:8.2f} us  {k}")
    json.dump({"model": a.model, "total_ms": total / 1e3, "rows": rows}, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
