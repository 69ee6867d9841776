"""BASELINE config 5: Robust Video Matting + DAMO text-to-video tasks mined together across the GPUs of the box
(one predictor per model per GPU, the scheduler packs the mixed queue), against the in-process chain."""
import json

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MINER, USER = "0x" + "a1" * 20, "0x" + "11" * 20


def test_mixed_rvm_and_damo_tasks_are_mined_across_the_gpus(tmp_path):
    from arbius_b200.chain.client import SimChainClient
    from arbius_b200.chain.engine_sim import UNIT, deploy_local
    from arbius_b200.codecs.mp4 import encode_mp4
    from arbius_b200.miner.db import MinerDB
    from arbius_b200.miner.ml import NativeBackend
    from arbius_b200.miner.node import MinerNode
    from arbius_b200.models.registry import build_predictor
    from arbius_b200.protocol import MiningFilter, ModelEntry, load_template, solution_cid, taskid2seed, template_bytes
    from arbius_b200.protocol.config import parse_mining_config
    n_gpus = 2 if torch.cuda.device_count() >= 2 else 1
    eng = deploy_local(engine_balance=599990 * UNIT, funded={MINER: 50 * UNIT, USER: 50 * UNIT})
    models, preds = [], {}
    for name in ("robust_video_matting", "damo"):
        mid = eng.register_model(USER, USER, 0, template_bytes(name))
        models.append(ModelEntry(id=mid, template=load_template(name), filters=[MiningFilter()], name=name))
        preds[mid] = [build_predictor(name, device=f"cuda:{g}", tiny=True) for g in range(n_gpus)]
    clip = (np.arange(3 * 64 * 96 * 3, dtype=np.uint32) * 2654435761 >> 24).astype(np.uint8).reshape(3, 64, 96, 3)
    video = tmp_path / "in.mp4"
    video.write_bytes(encode_mp4(clip, 8))
    cfg = parse_mining_config({"db_path": ":memory:", "ml": {"strategy": "native"}, "ipfs": {"strategy": "local"},
                               "automine": {"enabled": False}})
    node = MinerNode(cfg, SimChainClient(eng, MINER), NativeBackend(preds), models, db=MinerDB(":memory:"), n_gpus=n_gpus,
                     sleep=lambda s: None)
    node.bootup()
    node.run_until_idle()
    inputs = []
    for i in range(3):
        inputs.append((models[0], {"input_video": str(video), "output_type": "alpha-mask" if i % 2 else "green-screen"}))
        inputs.append((models[1], {"prompt": f"a panda eating bamboo, take {i}", "num_frames": 8, "num_inference_steps": 2}))
    tids = []
    for m, inp in inputs:
        eng.token.approve(USER, eng.address, 0)
        tids.append(eng.submit_task(USER, 0, USER, m.id, 0, json.dumps(inp, separators=(",", ":")).encode()))
    node.run_until_idle()
    assert node.db.get_failed_jobs() == []
    from arbius_b200.protocol import hydrate_input
    for (m, inp), tid in zip(inputs, tids):
        assert tid in eng.solutions, (m.name, inp)
        hyd = dict(hydrate_input(inp, m.template).input, seed=taskid2seed(tid))
        want = solution_cid({"out-1.mp4": preds[m.id][0](hyd)})          # the same task run directly on GPU 0
        assert eng.solutions[tid].cid == want, m.name
    used = {s["gpus"] for s in node.solved}
    assert len(node.solved) == 6 and (n_gpus == 1 or used == {(0,), (1,)})
