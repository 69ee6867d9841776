"""The miner's batching predictor on a real GPU: concurrent solves merged into one graph replay keep their bytes."""
import threading

import pytest

pytestmark = pytest.mark.gpu


def test_batching_predictor_merges_concurrent_solves_and_keeps_bytes():
    from arbius_b200.models.registry import BatchingPredictor, build_predictor
    single = build_predictor("anythingv3", device="cuda:0", tiny=True)
    batched = build_predictor("anythingv3", device="cuda:0", tiny=True, task_batch=4)
    assert isinstance(batched, BatchingPredictor)
    batched.window_s = 0.5                                   # the threads below start within this window
    inputs = [{"prompt": f"prompt {i}", "negative_prompt": "", "width": 64, "height": 64, "num_inference_steps": 2,
               "guidance_scale": 7, "scheduler": "DPMSolverMultistep", "seed": 40 + i} for i in range(6)]
    want = [single(i) for i in inputs]
    batched(inputs[0])                                       # warm-up: builds the single-task graph
    got = [None] * len(inputs)
    ths = [threading.Thread(target=lambda k=k: got.__setitem__(k, batched(inputs[k]))) for k in range(len(inputs))]
    for t in ths:
        t.start()
    for t in ths:
        t.join(120)
    assert got == want
    assert max(batched.batches) >= 2, batched.batches        # at least one group really was merged
    batched.close()
