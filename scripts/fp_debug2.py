"""Debug driver: the miner's predictor set in one process, then one frame-parallel solve from a worker thread."""
import faulthandler
import sys
import threading
import time

sys.path.insert(0, ".")
from arbius_b200.models.registry import (build_frame_parallel_predictor, build_pair_predictor,  # noqa: E402
                                         build_predictor)

what = sys.argv[1] if len(sys.argv) > 1 else "all"
inp = {"prompt": "a panda eating bamboo", "negative_prompt": "noisy", "num_frames": 8, "num_inference_steps": 2,
       "width": 256, "height": 256, "fps": 8, "guidance_scale": 17.5, "seed": 5}
keep = []
if what in ("all", "singles"):
    keep += [build_predictor("anythingv3", device=f"cuda:{g}", tiny=True, task_batch=2) for g in range(2)]
    keep += [build_predictor("zeroscopev2xl", device=f"cuda:{g}", tiny=True) for g in range(2)]
if what in ("all", "pairs"):
    keep += [build_pair_predictor("anythingv3", devices=(0, 1), tiny=True)]
    keep += [build_pair_predictor("zeroscopev2xl", devices=(0, 1), tiny=True)]
fp = build_frame_parallel_predictor("zeroscopev2xl", devices=(0, 1), tiny=True)
print("built", what, flush=True)
faulthandler.dump_traceback_later(6, exit=False)
res = {}
t = threading.Thread(target=lambda: res.__setitem__("out", fp(dict(inp))))
t0 = time.time()
t.start()
t.join(60)
faulthandler.cancel_dump_traceback_later()
print("done", what, round(time.time() - t0, 2), type(res.get("out")), flush=True)
