"""Tile/split-K table rules (CPU): split-K is part of the canonical summation order and must never depend on how many
tasks share a batch or how many GPUs share a task; the tile width may."""
import json
import sys
from fractions import Fraction
from pathlib import Path

import arbius_b200.ops  # noqa: F401

G = sys.modules["arbius_b200.ops.gemm"]


def test_split_k_always_comes_from_the_canonical_shape():
    table = json.loads(Path(G.__file__).with_name("tuned_sm100.json").read_text())
    assert table["conv:2x8x8:1280:1280"][1] > 1                     # a split-K shape of the lone task
    for n in (2, 4, 8):
        with G.batch_scale(Fraction(1, n)):
            assert G._canon(2 * n) == 2
            bn, sk = G._tuned_pair("conv:2x8x8:1280:1280", f"conv:{2 * n}x8x8:1280:1280")
            assert sk == table["conv:2x8x8:1280:1280"][1]
            own = table.get(f"conv:{2 * n}x8x8:1280:1280")
            assert bn == (own[0] if own else 0)                      # real shape's tile width, or the library's choice
    with G.batch_scale(2):                                           # guidance / frame split: canonical entry as is
        assert G._tuned_pair("conv:2x8x8:1280:1280", "conv:1x8x8:1280:1280") == table["conv:2x8x8:1280:1280"]
    assert G.get_batch_scale() == 1
    assert G._tuned_pair("conv:2x8x8:1280:1280", "conv:2x8x8:1280:1280") == table["conv:2x8x8:1280:1280"]
    assert G._tuned_pair("conv:3x5x7:64:64", "conv:3x5x7:64:64") is None


def test_batched_entries_keep_the_lone_tasks_split_k():
    """Every entry that exists for a 4-task batch shape carries the split-K of the shape it was derived from."""
    table = json.loads(Path(G.__file__).with_name("tuned_sm100.json").read_text())
    checked = 0
    for key, (bn, sk) in table.items():
        parts = key.split(":")
        if parts[0] == "conv" and parts[1].startswith("2x"):
            big = f"conv:8x{parts[1][2:]}:{parts[2]}:{parts[3]}"
            if big in table:
                assert table[big][1] == sk, (key, big)
                checked += 1
    assert checked >= 10
