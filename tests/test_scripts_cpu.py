"""The analysis scripts must keep working on the artefacts committed under profiles/."""

from __future__ import annotations

import py_compile
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_roofline_table_parses_a_committed_launch_list() -> None:
    csv = ROOT / "profiles" / "launches_step_r1_fused_bias.csv"
    proc = subprocess.run([sys.executable, str(ROOT / "scripts" / "roofline_table.py"), str(csv)],
                          capture_output=True, text=True, timeout=120, check=False)
    assert proc.returncode == 0, proc.stderr[-1500:]
    out = proc.stdout
    assert "gemm::gemm_kernel<1, 1, 4, 2>" in out and "TFLOP/s" in out
    assert "attn_bwd::bwd_kernel" in out and "optim::adamw_kernel" in out


@pytest.mark.parametrize(
    "path",
    sorted([*(ROOT / "bench").glob("*.py"), *(ROOT / "scripts").glob("*.py"), ROOT / "bench.py",
            ROOT / "__graft_entry__.py", ROOT / "baseline" / "ref_harness.py"]),
    ids=lambda p: p.name,
)
def test_tools_compile(path: Path, tmp_path: Path) -> None:
    py_compile.compile(str(path), cfile=str(tmp_path / "out.pyc"), doraise=True)
