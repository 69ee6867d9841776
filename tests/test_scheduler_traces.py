"""The planner's golden traces. ``miner/src/scheduler.ts`` (the scheduler for the unchanged TypeScript miner; reference:
miner/src/index.ts:879-958 awaits one solve at a time) mirrors ``arbius_b200/miner/scheduler.py`` function by function;
Node is absent on this box, so both implementations are pinned to ONE trace file: this test replays it against the
Python planner, ``miner/test/scheduler.test.ts`` replays it against the TypeScript one."""
import json
import re
from pathlib import Path

from arbius_b200.miner.scheduler import CostModel, SubsetPlanner, simulate

ROOT = Path(__file__).resolve().parents[1]
TRACES = json.loads((ROOT / "tests" / "data" / "scheduler_traces.json").read_text())


def test_python_planner_replays_the_golden_simulations():
    assert len(TRACES["simulations"]) >= 5
    for t in TRACES["simulations"]:
        cost = CostModel(table={f: {int(p): v for p, v in row.items()} for f, row in t["cost"].items()})
        r = simulate([tuple(x) for x in t["tasks"]], t["n_gpus"], cost, t["policy"])
        assert r["assignments"] == t["expect"]["assignments"], t["name"]
        assert abs(r["tasks_per_hour"] - t["expect"]["tasks_per_hour"]) <= 1e-6 * t["expect"]["tasks_per_hour"]
        assert abs(r["p50_latency_s"] - t["expect"]["p50_latency_s"]) <= 1e-9


def test_python_planner_replays_the_golden_planner_ops():
    for t in TRACES["planner"]:
        pl, got = SubsetPlanner(t["n_gpus"], t["slots"]), []
        for op in t["ops"]:
            if op[0] == "acquire":
                r = pl.acquire(op[1])
                got.append(list(r) if r else None)
            elif op[0] == "release":
                pl.release(op[1])
                got.append(None)
            else:
                got.append(pl.largest_free_aligned())
        assert got == t["expect"], t["name"]


def test_lone_video_gets_the_whole_box_and_a_burst_of_images_one_gpu_each():
    by = {t["name"]: t for t in TRACES["simulations"]}
    assert by["lone video, latency"]["expect"]["assignments"] == [[0, 1, 2, 3, 4, 5, 6, 7]]
    burst = by["burst of images, throughput"]["expect"]["assignments"]
    assert burst[:8] == [[g] for g in range(8)]             # while work is queued: one GPU per task, all eight busy


def test_typescript_sources_export_the_mirrored_api():
    """Structural check of the TS side (it cannot be executed here): the symbols the traces exercise exist, the config
    stays additive ({url} still accepted), solves get per-task cache directories."""
    sched = (ROOT / "miner" / "src" / "scheduler.ts").read_text()
    work = (ROOT / "miner" / "src" / "workers.ts").read_text()
    for sym in ("export class CostModel", "export class SubsetPlanner", "export function simulate", "bestThroughputP",
                "bestLatencyP", "largestFreeAligned", "chooseP"):
        assert sym in sched, sym
    for sym in ("export class WorkerRegistry", "export class SolvePump", "export async function solveOnWorker",
                "export function taskCacheDir", "pickWorker", "url?: string", "urls?: string[]"):
        assert sym in work, sym
    assert re.search(r"path\.join\(cacheRoot, taskid\)", work)            # cache/<taskid>/, not the shared out-1.png
    test_ts = (ROOT / "miner" / "test" / "scheduler.test.ts").read_text()
    assert "scheduler_traces.json" in test_ts
