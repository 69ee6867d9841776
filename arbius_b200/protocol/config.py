"""``MiningConfig.json`` loader — schema unchanged from the reference.

Reference: miner/src/types.ts:3-54 (shape), miner/MiningConfig.example.json
(canonical file incl. ``"//"`` comment keys), miner/src/start.ts:11-18 (argv[2],
plain JSON.parse, no validation). New keys are additive and optional so the
example file loads unchanged:

  ``ml.strategy: "native"``     run models in-process on the local B200s
  ``ml.native``                 {gpus, subsets, models:{modelid:{family, weights}}}
  ``ipfs.strategy: "local"``    compute the CID locally, pin best-effort
  ``scheduler``                 {policy, max_inflight, poll_ms}
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Optional


def _strip_comments(o: Any) -> Any:
    if isinstance(o, dict):
        return {k: _strip_comments(v) for k, v in o.items() if k != "//"}
    if isinstance(o, list):
        return [_strip_comments(v) for v in o]
    return o


@dataclass
class BlockchainConfig:
    private_key: str = ""
    rpc_url: str = ""
    use_delegated_validator: bool = False
    delegated_validator_address: str = ""


@dataclass
class RpcConfig:
    host: str = "localhost"
    port: int = 8335


@dataclass
class AutomineConfig:
    enabled: bool = False
    delay: int = 60
    version: int = 0
    model: str = ""
    fee: str = "0"
    input: dict = field(default_factory=dict)


@dataclass
class MlConfig:
    strategy: str = "cog"
    cog: dict = field(default_factory=dict)          # modelid -> {"url": ...}
    replicate: dict = field(default_factory=dict)
    native: dict = field(default_factory=dict)


@dataclass
class IpfsConfig:
    strategy: str = "http_client"
    http_client: dict = field(default_factory=lambda: {"url": "http://127.0.0.1:5001"})
    pinata: dict = field(default_factory=dict)


@dataclass
class SchedulerConfig:
    policy: str = "throughput"      # or "latency"
    max_inflight: int = 0           # 0 = one per GPU subset
    poll_ms: int = 100              # reference poll interval (miner/src/index.ts:1081)
    gpus: int = 0                   # 0 = autodetect


@dataclass
class MiningConfig:
    log_path: Optional[str] = None
    db_path: str = "db.sqlite"
    stake_buffer_percent: int = 20
    stake_buffer_topup_percent: int = 1
    evilmode: bool = False
    blockchain: BlockchainConfig = field(default_factory=BlockchainConfig)
    rpc: RpcConfig = field(default_factory=RpcConfig)
    automine: AutomineConfig = field(default_factory=AutomineConfig)
    ml: MlConfig = field(default_factory=MlConfig)
    ipfs: IpfsConfig = field(default_factory=IpfsConfig)
    scheduler: SchedulerConfig = field(default_factory=SchedulerConfig)
    raw: dict = field(default_factory=dict, repr=False)


def _take(cls, d: Optional[dict]):
    d = d or {}
    names = {f for f in cls.__dataclass_fields__}
    return cls(**{k: v for k, v in d.items() if k in names})


def parse_mining_config(obj: dict) -> MiningConfig:
    raw = obj
    o = _strip_comments(obj)
    return MiningConfig(
        log_path=o.get("log_path"),
        db_path=o.get("db_path", "db.sqlite"),
        stake_buffer_percent=o.get("stake_buffer_percent", 20),
        stake_buffer_topup_percent=o.get("stake_buffer_topup_percent", 1),
        evilmode=bool(o.get("evilmode", False)),
        blockchain=_take(BlockchainConfig, o.get("blockchain")),
        rpc=_take(RpcConfig, o.get("rpc")),
        automine=_take(AutomineConfig, o.get("automine")),
        ml=_take(MlConfig, o.get("ml")),
        ipfs=_take(IpfsConfig, o.get("ipfs")),
        scheduler=_take(SchedulerConfig, o.get("scheduler")),
        raw=raw,
    )


def load_mining_config(path: str | Path) -> MiningConfig:
    """Raises ValueError on unparsable JSON (the reference exits, start.ts:15-18)."""
    try:
        return parse_mining_config(json.loads(Path(path).read_text()))
    except json.JSONDecodeError as exc:  # pragma: no cover - message path
        raise ValueError(f"unable to parse {path}: {exc}") from exc


# Static chain config shipped by the reference (miner/src/config.json:1-18).
STATIC_CHAIN_CONFIG = {
    "l1TokenAddress": "0xe3DBC4F88EAa632DDF9708732E2832EEaA6688AB",
    "baseTokenAddress": "0xe3DBC4F88EAa632DDF9708732E2832EEaA6688AB",
    "engineAddress": "0x399511EDEB7ca4A8328E801b1B3D0fe232aBc996",
    "proxyAdminAddress": "0xf70a86d51Bd88885054ea031344447d75bb4b432",
    "models": {
        "kandinsky2": {
            "id": "0x82ae0d19f32b6912204160f8d744de08265b7775d07c23b4171c94c8e2892c62",
            "mineable": True,
            "contracts": {},
            "params": {
                "addr": "0x0000000000000000000000000000000000000001",
                "fee": "0",
                "rate": "1000000000000000000",
                "cid": "0x1220511fdf0e88fa9adba98a7693cca89b5d9f3181da4815b0f0500aa4315b38d17d",
            },
        }
    },
}
