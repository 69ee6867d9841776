"""bench.py contract pieces that do not need a GPU."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_prints_one_json_line_and_exits_zero():
    env = dict(os.environ, RANK="0")
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "3"], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0, r.stderr[-500:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and isinstance(d["unavailable"], str) and "\n" not in d["unavailable"]
    # other ranks of a torchrun launch stay silent
    r1 = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference"], capture_output=True, text=True,
                        timeout=120, env=dict(os.environ, RANK="1"))
    assert r1.returncode == 0 and r1.stdout.strip() == ""


def test_graft_entry_has_build_and_smoke():
    sys.path.insert(0, str(ROOT))
    import __graft_entry__ as g
    assert callable(g.build) and callable(g.smoke)
