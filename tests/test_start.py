"""`python -m arbius_b200.miner MiningConfig.json` boot path (reference: miner/src/start.ts:11-59), CPU only."""
import json

import pytest

from arbius_b200.miner import start as start_mod
from arbius_b200.miner.ml import CogBackend, encode_cog_output
from arbius_b200.protocol.config import STATIC_CHAIN_CONFIG


def _config(tmp_path, **over):
    kid = STATIC_CHAIN_CONFIG["models"]["kandinsky2"]["id"]
    cfg = {"log_path": str(tmp_path / "log.txt"), "db_path": str(tmp_path / "db.sqlite"), "stake_buffer_percent": 20,
           "stake_buffer_topup_percent": 1, "evilmode": False,
           "blockchain": {"private_key": "0x" + "11" * 32, "rpc_url": "sim://"},
           "rpc": {"host": "127.0.0.1", "port": 0},
           "ml": {"strategy": "cog", "cog": {kid: {"url": "http://127.0.0.1:1/predictions", "//": "comment keys stay legal"}}},
           "ipfs": {"strategy": "local"}, "automine": {"enabled": False}}
    cfg.update(over)
    p = tmp_path / "MiningConfig.json"
    p.write_text(json.dumps(cfg))
    return str(p)


def test_boot_builds_a_node_against_the_in_process_chain(tmp_path, capsys):
    node = start_mod.start(_config(tmp_path), run=False)
    assert isinstance(node.ml, CogBackend) and node.pool.planner.n == 1
    assert [m.name for m in node.models] == ["kandinsky2"]                 # the statically known id -> its template
    # the node can solve through a (faked) Cog endpoint end to end
    node.ml._post = lambda url, payload, timeout: encode_cog_output(b"\x89PNG fake", "image/png")
    node.bootup()
    node.run_until_idle()
    assert (tmp_path / "log.txt").exists() and (tmp_path / "db.sqlite").exists()
    assert "EVILMODE" not in capsys.readouterr().err


def test_bad_config_exits_1_and_evilmode_warns_loudly(tmp_path, capsys):
    bad = tmp_path / "bad.json"
    bad.write_text("{not json")
    with pytest.raises(SystemExit) as e:
        start_mod.start(str(bad), run=False)
    assert e.value.code == 1 and "unable to parse" in capsys.readouterr().err
    start_mod.start(_config(tmp_path, evilmode=True), run=False)
    assert capsys.readouterr().err.count("EVILMODE ENABLED") == 20         # start.ts:20-25 prints its warning in a loop
    assert start_mod.main([]) == 1                                         # usage
