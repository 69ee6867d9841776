"""Per-kernel breakdown of one RVM frame at 1080p (no graph), torch.profiler."""
import re, sys
from collections import defaultdict
import numpy as np, torch
sys.path.insert(0, ".")
from arbius_b200.models.rvm.pipeline import RvmPipeline
from arbius_b200.models.rvm.ref import RvmConfig
pipe = RvmPipeline(RvmConfig(), backend="native", device="cuda", dtype=torch.float16, use_graph=False)
clip = np.random.RandomState(0).randint(0, 255, (4, 1080, 1920, 3)).astype(np.uint8)
pipe.matte(clip, "green-screen")
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    pipe.matte(clip, "green-screen")
    torch.cuda.synchronize()
agg = defaultdict(lambda: [0.0, 0])
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        key = re.sub(r"<.*", "", ev.name)
        agg[key][0] += ev.device_time; agg[key][1] += 1
tot = sum(v[0] for v in agg.values())
print(f"total {tot/1e3/4:.3f} ms per frame")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:14]:
    print(f"{v[0]/1e3/4:8.3f} ms/frame {100*v[0]/tot:5.1f}% {v[1]//4:4d}x avg {v[0]/v[1]:7.1f} us  {k[:60]}")
