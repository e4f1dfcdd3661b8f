"""Real multi-GPU slab run (needs >= 2 GPUs; skipped otherwise): torchrun with one rank per GPU,
both transports (direct peer stores + handshake, NCCL send/recv + unpack) against the undivided
run."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("stencil,collision", [("D3Q19", "bgk"), ("D3Q27", "kbc")])
def test_slab_run_matches_undivided(stencil, collision):
    n = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29517",
           os.path.join(ROOT, "scripts", "slab_check.py"), "--res", "48", "--stencil", stencil,
           "--collision", collision]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "slab_check OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
