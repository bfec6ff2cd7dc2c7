"""NVLink peer-memory communicator: needs >= 2 GPUs on the box (skipped otherwise)."""

from __future__ import annotations

import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return int(s.getsockname()[1])


@pytest.mark.parametrize("world", [2, 4, 8])
def test_nvl_allreduce_step_and_trainer(world: int) -> None:
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    worker = Path(__file__).with_name("_nvl_worker.py")
    proc = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
         "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(worker)],
        capture_output=True, text=True, timeout=600, check=False,
    )
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    assert "NVL_WORKER_OK" in proc.stdout
    assert proc.stdout.count("DDP_EQUIVALENCE_OK") == 3, proc.stdout[-2000:]


def test_nvls_multimem_allreduce() -> None:
    """In-switch reduction (multimem.ld_reduce / multimem.st); skipped where multicast is absent."""
    world = 2
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    worker = Path(__file__).with_name("_nvl_worker.py")
    proc = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
         "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(worker), "nvls"],
        capture_output=True, text=True, timeout=600, check=False,
    )
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    assert "NVL_WORKER_OK" in proc.stdout
    if "NVLS_UNAVAILABLE" in proc.stdout:
        pytest.skip("multicast objects cannot be created / shared on this box")
    assert "NVLS_OK" in proc.stdout, proc.stdout[-2000:]
