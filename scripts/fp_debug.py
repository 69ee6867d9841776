"""Debug driver: in-process frame-parallel predictor at a given size (python scripts/fp_debug.py W H)."""
import sys
import time

sys.path.insert(0, ".")
from arbius_b200.models.registry import build_frame_parallel_predictor, build_predictor  # noqa: E402

w, h = int(sys.argv[1]), int(sys.argv[2])
inp = {"prompt": "a panda eating bamboo", "negative_prompt": "noisy", "num_frames": 8, "num_inference_steps": 2,
       "width": w, "height": h, "fps": 8, "guidance_scale": 17.5, "seed": 5}
single = build_predictor("zeroscopev2xl", device="cuda:0", tiny=True)
t0 = time.time()
want = single(dict(inp))
print("single ok", round(time.time() - t0, 2), flush=True)
fp = build_frame_parallel_predictor("zeroscopev2xl", devices=(0, 1), tiny=True)
t0 = time.time()
got = fp(dict(inp))
print("frame-parallel ok", round(time.time() - t0, 2), got == want, flush=True)
print("replay", fp(dict(inp)) == want, flush=True)
