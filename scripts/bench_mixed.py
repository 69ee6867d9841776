"""BASELINE config 5 at real sizes: Robust Video Matting 1080p clips + DAMO text-to-video (256x256x16, 50 steps) mined as
one mixed queue over the GPUs of a box (torchrun, one process per GPU = the scheduler's throughput policy for
single-GPU tasks: the i-th task of the queue goes to GPU i mod N). Every task goes through the public ``predict`` call
(decode / H2D in, MP4 bytes out). Device-timed between barriers, max over ranks; one JSON line with tasks/hour per
family.

  python -m torch.distributed.run --nproc-per-node 8 scripts/bench_mixed.py --tasks-per-gpu 4
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tasks-per-gpu", type=int, default=4, help="tasks per GPU in the timed region (alternating families)")
    ap.add_argument("--frames", type=int, default=48, help="frames per RVM clip (1080p)")
    a = ap.parse_args()
    from arbius_b200.models.rvm.pipeline import RvmPipeline
    from arbius_b200.models.rvm.ref import RvmConfig
    from arbius_b200.models.video.pipeline import TextToVideoConfig, TextToVideoPipeline
    from arbius_b200.parallel import init_distributed
    from arbius_b200.protocol import solution_cid
    world = init_distributed()
    rank = dist.get_rank() if world > 1 else 0
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    rvm = RvmPipeline(RvmConfig(), backend="native", device=dev, dtype=torch.float16)
    damo = TextToVideoPipeline(TextToVideoConfig.damo(), backend="native", device=dev, dtype=torch.float16, family="damo")
    rng = np.random.RandomState(rank)
    clip = rng.randint(0, 255, (a.frames, 1080, 1920, 3)).astype(np.uint8)
    dinp = {"prompt": "a panda eating bamboo on a rock", "num_frames": 16, "num_inference_steps": 50, "fps": 8}

    def run(i):
        """Task i of this rank: the two families alternate over the queue and are shifted by the rank, so at any moment half
        of the GPUs matte a clip while the other half denoise a video."""
        if (i + rank) % 2 == 0:
            return "robust_video_matting", rvm.predict({"input_video": clip, "output_type": "green-screen", "seed": i})
        return "damo", damo.predict(dict(dinp, seed=1000 * rank + i))
    for i in range(4):                                            # warm-up: both families' graphs get captured, twice each
        run(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    counts = {"robust_video_matting": 0, "damo": 0}
    cid = None
    e0.record()
    for i in range(a.tasks_per_gpu):
        fam, blob = run(i)
        counts[fam] += 1
        cid = solution_cid({"out-1.mp4": blob})
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms, counts["robust_video_matting"], counts["damo"]], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms = mx[0].item()
    if rank == 0:
        n_rvm, n_damo = int(t[1].item()), int(t[2].item())
        print(json.dumps({"bench": "mixed mining: RVM 1080p clips + DAMO text-to-video 256x256x16/50 steps", "n_gpus": world,
                          "timed_ms": round(ms, 1), "tasks": {"robust_video_matting": n_rvm, "damo": n_damo},
                          "tasks_per_hour": {"robust_video_matting": round(n_rvm / ms * 3.6e6, 1), "damo": round(n_damo / ms * 3.6e6, 1),
                                             "total": round((n_rvm + n_damo) / ms * 3.6e6, 1)},
                          "rvm_frames_per_clip": a.frames, "rvm_frames_per_second_box": round(n_rvm * a.frames / ms * 1e3, 1),
                          "includes": "per task: decode/H2D, model, D2H, MP4 encode, CIDv0", "last_cid": cid}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
