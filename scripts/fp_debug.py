"""In-process frame-parallel predictor over N GPUs of this box vs one GPU (python scripts/fp_debug.py W H FRAMES NGPU)."""
import sys
import time

sys.path.insert(0, ".")
from arbius_b200.models.registry import build_frame_parallel_predictor, build_predictor  # noqa: E402

w, h, f, n = (int(v) for v in (sys.argv[1:5] + ["64", "64", "8", "2"][len(sys.argv) - 1:]))
inp = {"prompt": "a panda eating bamboo", "negative_prompt": "noisy", "num_frames": f, "num_inference_steps": 2,
       "width": w, "height": h, "fps": 8, "guidance_scale": 17.5, "seed": 5}
single = build_predictor("zeroscopev2xl", device="cuda:0", tiny=True)
want = single(dict(inp))
fp = build_frame_parallel_predictor("zeroscopev2xl", devices=tuple(range(n)), tiny=True)
t0 = time.time()
got = fp(dict(inp))
print(f"frame-parallel x{n} {w}x{h}x{f}: first call {time.time() - t0:.2f} s, bytes equal: {got == want}", flush=True)
t0 = time.time()
print(f"replay bytes equal: {fp(dict(inp)) == want} ({time.time() - t0:.3f} s)", flush=True)
