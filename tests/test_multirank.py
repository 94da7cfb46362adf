"""world_size = 2 coverage of the row-sharded path: gloo on CPU for the exchange
structure, NCCL on two B200s (when the box has them) for the real kernels + dr4g_comm_init."""
from __future__ import annotations

import socket
import subprocess
import sys
from pathlib import Path

import pytest

WORKER = Path(__file__).resolve().parent / "multirank_worker.py"


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch(mode: str, n: int, d: int, k: int, timeout: int) -> str:
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(WORKER), mode, str(n), str(d), str(k)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def test_two_ranks_gloo_packed_allreduce_matches_oracle():
    out = _launch("cpu", 3000, 48, 6, 300)
    assert "MULTIRANK_CPU_OK" in out


@pytest.mark.gpu
def test_two_gpus_nccl_sharded_fit_matches_reference():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on the box")
    out = _launch("gpu", 40000, 256, 16, 240)
    assert "MULTIRANK_GPU_OK" in out


@pytest.mark.gpu
def test_two_gpus_row_folded_narrow_shards_match_reference():
    """d = 32 shards of 600,000 rows: each rank's pass is row-folded 8-fold; the sharded model against the reference backend."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on the box")
    out = _launch("gpu-fold", 600000, 32, 8, 300)
    assert "MULTIRANK_FOLD_OK" in out
