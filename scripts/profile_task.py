"""Per-kernel time breakdown of one AnythingV3 task (CUPTI via torch.profiler; shares, not absolutes)."""
import argparse
import json
import re
import sys
from collections import defaultdict

import torch

sys.path.insert(0, ".")
from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="native")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--out", default="gpurun_out/profile_task.json")
    a = ap.parse_args()
    pipe = SD15Pipeline(SD15Config(), backend=a.impl, device="cuda", dtype=torch.float16, use_graph=False)
    kw = dict(prompt="arbius test cat", negative_prompt="", width=a.size, height=a.size,
              num_inference_steps=a.steps, guidance_scale=12.0, scheduler="DPMSolverMultistep")
    pipe.generate(seed=1, **kw)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        pipe.generate(seed=2, **kw)
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
    print(f"total kernel time {total / 1e3:.2f} ms over {sum(r[2] for r in rows)} launches")
    for k, ms, n in rows[:25]:
        print(f"{ms:9.3f} ms {100 * ms * 1e3 / total:5.1f}% {n:6d}x  avg {1e3 * ms / n:8.2f} us  {k}")
    json.dump({"total_ms": total / 1e3, "rows": rows}, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
