"""Multi-GPU tests: each spawns torchrun workers (one process per GPU) and checks their exit status."""
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def _torchrun(script, nproc, port, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, script)]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)


@pytest.mark.gpu
@pytest.mark.multigpu
def test_patch_parallel_attention_is_bit_identical_to_single_gpu():
    n = min(torch.cuda.device_count(), 8)
    n = max(p for p in (2, 4, 8) if p <= n)
    r = _torchrun("mp_peer_attention.py", n, 29611)
    sys.stdout.write(r.stdout[-3000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "bit-identical=False" not in r.stdout and "RESULT" in r.stdout


@pytest.mark.gpu
@pytest.mark.multigpu
def test_fused_tp_gemms_are_bit_identical_to_single_gpu():
    n = min(torch.cuda.device_count(), 8)
    n = max(p for p in (2, 4, 8) if p <= n)
    r = _torchrun("mp_tp_gemm.py", n, 29612)
    sys.stdout.write(r.stdout[-3000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "bit-identical=False" not in r.stdout and "RESULT" in r.stdout
