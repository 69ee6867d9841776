"""Per-CTA phase timing (scripts/gemm_phases.py probes) of the large-M projections of ZeroScope XL under full load.
AB_GEMM_PERSIST=0 shows the one-tile-per-CTA launch."""
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "scripts")
import gemm_phases as gp  # noqa: E402
from arbius_b200 import ops  # noqa: E402

torch.manual_seed(0)
for (m, n, k, bn) in [(442368, 320, 320, 160), (110592, 640, 640, 160), (442368, 320, 1280, 160)]:
    x = torch.randn(m, k, device="cuda").half()
    w = (torch.randn(n, k, device="cuda") * 0.02).half()
    b = torch.randn(n, device="cuda").half()
    gp.run(f"linear {m}x{n}x{k} bn{bn}", lambda: ops.gemm(x, w, b, block_n=bn, splitk=1), grid_hint=8192)
