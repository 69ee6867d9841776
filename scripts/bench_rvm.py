"""RVM frame throughput: native engine (CUDA graph per frame) vs the eager fp16 oracle, 1080p, downsample_ratio 0.25.
Device-timed with CUDA events after warm-up; prints one JSON line per backend."""
import argparse
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from arbius_b200.models.rvm.pipeline import RvmPipeline  # noqa: E402
from arbius_b200.models.rvm.ref import RvmConfig  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=48)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    rng = np.random.RandomState(0)
    clip = rng.randint(0, 255, (a.frames, a.height, a.width, 3)).astype(np.uint8)
    for backend, dtype in (("native", torch.float16), ("eager", torch.float16)):
        pipe = RvmPipeline(RvmConfig(), backend=backend, device="cuda", dtype=dtype)
        pipe.matte(clip[:4], "green-screen")
        pipe.matte(clip[:4], "green-screen")
        times = []
        for _ in range(a.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            pipe.matte(clip, "green-screen")
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        ms = min(times) / a.frames
        print(json.dumps({"bench": "rvm_1080p", "backend": backend, "ms_per_frame_e2e": round(ms, 3),
                          "fps": round(1000.0 / ms, 1), "frames": a.frames,
                          "launches_per_frame": getattr(pipe, "last_launches", 0) // max(1, a.frames)}), flush=True)


if __name__ == "__main__":
    main()
