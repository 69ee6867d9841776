"""The miner's native boot path on a multi-GPU box: MiningConfig -> predictors per GPU, pair / frame-parallel subset
predictors, cost table -> a lone task is widened over idle GPUs (latency policy) and its CID equals the 1-GPU one."""
import json

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def test_native_config_builds_subset_predictors_and_widens_a_lone_video_task(tmp_path):
    from arbius_b200.miner import start as start_mod
    from arbius_b200.models.registry import build_predictor
    from arbius_b200.protocol import hydrate_input, load_template, solution_cid, taskid2seed
    from arbius_b200.protocol.config import parse_mining_config
    _, registered = start_mod._sim_chain(parse_mining_config({}))          # model ids are deterministic on the sim chain
    ids = {name: mid for mid, name in registered.items()}
    cfg = {"db_path": str(tmp_path / "db.sqlite"), "blockchain": {"private_key": "0x" + "11" * 32, "rpc_url": "sim://"},
           "rpc": {"host": "127.0.0.1", "port": 0}, "ipfs": {"strategy": "local"}, "automine": {"enabled": False},
           "scheduler": {"policy": "latency"},
           "ml": {"strategy": "native", "native": {"gpus": 2, "pairs": True, "frame_parallel": [2], "task_batch": 2,
                  "models": {ids["zeroscopev2xl"]: {"template": "zeroscopev2xl", "tiny": True},
                             ids["anythingv3"]: {"template": "anythingv3", "tiny": True}}}}}
    path = tmp_path / "MiningConfig.json"
    path.write_text(json.dumps(cfg))
    node = start_mod.start(str(path), run=False)
    assert node.pool.planner.n == 2 and node.pool.planner.slots == 2
    assert set(node.ml.subset_sizes(next(m for m in node.models if m.name == "zeroscopev2xl"))) == {1, 2}
    eng = node.chain.engine
    node._retry_sleep = lambda s: None                                     # a failure should fail fast here
    node.bootup()
    node.run_until_idle()
    inp = {"prompt": "a panda eating bamboo", "negative_prompt": "noisy", "num_frames": 8, "num_inference_steps": 2,
           "width": 256, "height": 256}
    mid = ids["zeroscopev2xl"]
    tid = eng.submit_task(node.chain.address, 0, node.chain.address, mid, 0, json.dumps(inp, separators=(",", ":")).encode())
    node.run_until_idle()
    assert node.db.get_failed_jobs() == [] and tid in eng.solutions
    assert node.solved[-1]["gpus"] == (0, 1)                               # the lone clip took both GPUs
    hyd = dict(hydrate_input(inp, load_template("zeroscopev2xl")).input, seed=taskid2seed(tid))
    single = build_predictor("zeroscopev2xl", device="cuda:0", tiny=True)
    assert eng.solutions[tid].cid == solution_cid({"out-1.mp4": single(hyd)})
    # two image tasks at once: one per GPU or merged — either way both are solved with the lone-task bytes
    sd = {"prompt": "arbius test cat", "negative_prompt": "", "width": 128, "height": 128, "num_inference_steps": 2}
    t2 = [eng.submit_task(node.chain.address, 0, node.chain.address, ids["anythingv3"], 0,
                          json.dumps(dict(sd, prompt=f"cat {i}"), separators=(",", ":")).encode()) for i in range(3)]
    node.run_until_idle()
    assert all(t in eng.solutions for t in t2) and node.db.get_failed_jobs() == []
