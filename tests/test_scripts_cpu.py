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


def test_both_bench_arms_emit_the_same_config_object() -> None:
    """The driver compares the two arms' ``config`` dictionaries: both come from one function, and
    implementation-specific descriptions live under ``notes``."""
    import inspect

    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "baseline"))
    import bench
    import ref_harness

    ours = bench.bench_config(32, 1, 1024, 8)
    assert ours == bench.bench_config(32, 1, 1024, 8, "small")
    assert ours["model"].startswith("gpt2-small-124M") and ours["parallelism"] == "dp8"
    assert "comm" not in ours and "reference_mode" not in ours
    assert "config" in inspect.signature(ref_harness.run).parameters
    src = (ROOT / "baseline" / "ref_harness.py").read_text()
    assert '"config": config' in src and '"notes"' in src
    assert bench.bench_config(4, 1, 2048, 8, "xl")["model"].startswith("gpt2-xl")


def test_device_resident_tokens_need_a_cuda_run() -> None:
    from llmtrain.config.schemas import RunConfig
    from llmtrain.data.synthetic_text import SyntheticTextDataModule

    cfg = RunConfig.model_validate({
        "run": {"name": "t", "device": "cpu"},
        "model": {"name": "gpt", "block_size": 16, "d_model": 64, "n_layers": 1, "n_heads": 1,
                  "d_ff": 64, "vocab_size": 100},
        "data": {"name": "synthetic_text", "extra": {"device_resident": True}},
        "trainer": {"max_steps": 1, "micro_batch_size": 2, "warmup_steps": 0},
        "ddp": {}, "mlflow": {"enabled": False}, "logging": {}, "output": {},
    })
    module = SyntheticTextDataModule()
    module.setup(cfg)
    with pytest.raises(ValueError, match="device_resident"):
        module.train_dataloader()
