"""GPU tests of the data-parallel exchange step: the peer-memory all-reduce kernel with every rank played by one GPU
(tests/peer_check.py), and -- where the box has two or more GPUs -- tests/dp_check.py under torchrun with both
transports (peer memory, NCCL). Each runs in a subprocess with a time limit."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _run(cmd, env=None, limit=300):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run(cmd, cwd=ROOT, env=e, capture_output=True, text=True, timeout=limit)
    assert p.returncode == 0, f"{' '.join(cmd)} failed:\n{p.stdout[-3000:]}\n{p.stderr[-3000:]}"
    return p.stdout


def test_peer_allreduce_kernel_all_ranks_on_one_gpu():
    assert "peer_check ok" in _run([sys.executable, "tests/peer_check.py"])


@pytest.mark.parametrize("transport", ["peer", "nccl"])
def test_data_parallel_step_on_two_gpus(transport):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                "127.0.0.1", "--master-port", "29541", "tests/dp_check.py"], env={"B200VIT_DP_ALLREDUCE": transport})
    assert f"dp_check ok: world=2" in out and f"transport={transport}" in out
