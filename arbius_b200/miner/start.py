"""Miner entry point: ``python -m arbius_b200.miner MiningConfig.json``.

Boot order follows the reference (reference: miner/src/start.ts:11-59): parse config (exit on bad JSON),
evilmode banner, logger, git revision, database, ML backend, blockchain client, operator RPC, main loop.
Which chain transport is used comes from ``blockchain.rpc_url``: ``sim://`` starts an in-process EngineSim
with a bootstrap model (the offline stand-in for ``npx hardhat node`` + deploy scripts), anything else is
treated as an Ethereum JSON-RPC endpoint.
"""
from __future__ import annotations

import subprocess
import sys
from typing import List, Optional

from ..protocol.config import STATIC_CHAIN_CONFIG, MiningConfig, load_mining_config
from ..protocol.hydrate import MiningFilter, ModelEntry
from ..protocol.templates import TEMPLATE_NAMES, load_template, template_bytes
from .db import MinerDB
from .ml import make_backend
from .node import MinerNode
from .rpc import initialize_rpc
from .utils import initialize_logger, log


def enabled_models(c: MiningConfig, registered: Optional[dict] = None) -> List[ModelEntry]:
    """The reference hard-codes Kandinsky2 (index.ts:781-877); here every template can be enabled:
    ``ml.cog`` / ``ml.native.models`` keys are model ids; a template name may be given per id (default: the
    id registered for Kandinsky2 in the static chain config)."""
    known = {STATIC_CHAIN_CONFIG["models"]["kandinsky2"]["id"].lower(): "kandinsky2"}
    for mid, name in (registered or {}).items():
        known[mid.lower()] = name
    out = []
    ids = set(k.lower() for k in (c.ml.cog or {})) | set(k.lower() for k in (c.ml.native.get("models", {}) if c.ml.native else {}))
    for mid in sorted(ids):
        ent = (c.ml.native.get("models", {}) if c.ml.native else {}).get(mid) or (c.ml.cog or {}).get(mid) or {}
        name = ent.get("template") or known.get(mid)
        if name not in TEMPLATE_NAMES:
            log.warning("model %s has no known template; skipping", mid)
            continue
        out.append(ModelEntry(id=mid, template=load_template(name), mineable=True,
                              filters=[MiningFilter(minfee=int(ent.get("minfee", 0)), mintime=int(ent.get("mintime", 0)))],
                              name=name))
    return out


def _sim_chain(c: MiningConfig):
    from ..chain.client import SimChainClient
    from ..chain.engine_sim import deploy_local
    wallet = "0x" + "a1" * 20
    eng = deploy_local(funded={wallet: 100 * 10 ** 18})
    registered = {}
    for name in TEMPLATE_NAMES:
        mid = eng.register_model(wallet, "0x" + "00" * 19 + "01", 0, template_bytes(name))
        registered[mid] = name
    return SimChainClient(eng, wallet), registered


def start(config_path: str, run: bool = True) -> MinerNode:
    try:
        c = load_mining_config(config_path)
    except (OSError, ValueError) as e:
        print(f"unable to parse {config_path}: {e}", file=sys.stderr)
        raise SystemExit(1)
    if c.evilmode:
        for _ in range(20):
            print("WARNING: EVILMODE ENABLED — solutions will be wrong on purpose; only use on testnets", file=sys.stderr)
    initialize_logger(c.log_path)
    try:
        rev = subprocess.run(["git", "rev-parse", "HEAD"], capture_output=True, text=True, timeout=5).stdout.strip()
        log.info("Arbius miner %s starting", rev or "unknown")
    except (OSError, subprocess.SubprocessError):
        log.info("Arbius miner starting")
    db = MinerDB(c.db_path)
    registered = {}
    if c.blockchain.rpc_url.startswith("sim://") or c.blockchain.rpc_url in ("", "CHANGEME"):
        chain, registered = _sim_chain(c)
    else:
        from ..chain.jsonrpc import JsonRpcChainClient
        chain = JsonRpcChainClient(c.blockchain.rpc_url, c.blockchain.private_key,
                                   STATIC_CHAIN_CONFIG["engineAddress"], STATIC_CHAIN_CONFIG["baseTokenAddress"])
    models = enabled_models(c, registered)
    predictors = {}
    # ``ml.native.task_batch: n`` lets n solves share a GPU: the SD-1.5 predictor merges them into one graph replay
    # (more tasks/hour, same bytes per task); the pool and the backend then admit n solves per GPU.
    slots = int(c.ml.native.get("task_batch", 1)) if (c.ml.strategy == "native" and c.ml.native) else 1
    if c.ml.strategy == "native":
        from ..models.registry import build_predictor
        gpus = int(c.ml.native.get("gpus", 1))
        for m in models:
            ent = c.ml.native.get("models", {}).get(m.id, {})
            predictors[m.id] = [build_predictor(m.name, device=f"cuda:{g}", model_dir=ent.get("weights"),
                                                tiny=bool(ent.get("tiny")), task_batch=slots) for g in range(gpus)]
    # ``ml.native.pairs: true`` also builds one guidance-parallel predictor per aligned GPU pair for the families that
    # have one; the cost model then knows their measured 2-GPU latency and the latency policy may use them.
    subset_predictors, pair_families, cost = {}, [], None
    if c.ml.strategy == "native":
        from ..models.registry import PAIR_FAMILIES, build_pair_predictor
        from .scheduler import CostModel
        gpus = int(c.ml.native.get("gpus", 1))
        if c.ml.native.get("pairs") and gpus >= 2:
            for m in models:
                if m.name in PAIR_FAMILIES:
                    ent = c.ml.native.get("models", {}).get(m.id, {})
                    subset_predictors[m.id] = {(g, g + 1): build_pair_predictor(m.name, devices=(g, g + 1),
                                                                                model_dir=ent.get("weights"),
                                                                                tiny=bool(ent.get("tiny")))
                                               for g in range(0, gpus - 1, 2)}
                    pair_families.append(m.name)
        # ``ml.native.frame_parallel: [2, 4, 8]`` (or true = the whole box): text-to-video tasks may be spread over
        # aligned GPU subsets of those sizes, frames sharded (models/registry.py build_frame_parallel_predictor)
        fp_sizes = c.ml.native.get("frame_parallel") or []
        if fp_sizes is True:
            fp_sizes = [gpus]
        frame_families = []
        for m in models:
            from ..models.registry import FRAME_FAMILIES, build_frame_parallel_predictor
            if m.name not in FRAME_FAMILIES:
                continue
            ent = c.ml.native.get("models", {}).get(m.id, {})
            for p in sorted(int(x) for x in fp_sizes):
                if p not in (2, 4, 8) or p > gpus:
                    continue
                for g in range(0, gpus - p + 1, p):
                    devs = tuple(range(g, g + p))
                    subset_predictors.setdefault(m.id, {})[devs] = build_frame_parallel_predictor(
                        m.name, devices=devs, model_dir=ent.get("weights"), tiny=bool(ent.get("tiny")))
                if m.name not in frame_families:
                    frame_families.append(m.name)
        cost = CostModel.measured_b200(pair_families, frame_families, [int(x) for x in fp_sizes])
    ml = make_backend(c, predictors, subset_predictors, slots)
    n_gpus = c.scheduler.gpus or max([ml.endpoints(m) for m in models] + [1])
    n_gpus = max(p for p in (1, 2, 4, 8) if p <= n_gpus)
    from .golden import GoldenStore, default_path
    node = MinerNode(c, chain, ml, models, db=db, n_gpus=n_gpus, cost=cost, pin=c.ipfs.strategy != "local",
                     slots=slots, goldens=GoldenStore(default_path(c)))
    initialize_rpc(db, c.rpc.host, c.rpc.port, node)
    if run:
        node.run_forever()
    return node


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) < 1:
        print("usage: python -m arbius_b200.miner MiningConfig.json", file=sys.stderr)
        return 1
    start(argv[0])
    return 0
