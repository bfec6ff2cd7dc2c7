"""Process-group bootstrap and the data-parallel branches of the trainer, on CPU / gloo."""

from __future__ import annotations

import json
import os
import socket
import subprocess
import sys
from collections.abc import Iterator
from pathlib import Path
from unittest.mock import Mock

import pytest
import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel

from llmtrain.distributed import DDPState, setup_ddp, teardown_ddp
from llmtrain.training import Trainer
from tests.helpers import config, payload, write_yaml

SRC = str(Path(__file__).resolve().parents[1] / "src")
_VARS = ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return int(s.getsockname()[1])


@pytest.fixture(autouse=True)
def _clean_group(monkeypatch: pytest.MonkeyPatch) -> Iterator[None]:
    for var in _VARS:
        monkeypatch.delenv(var, raising=False)
    yield
    teardown_ddp()


def _single_rank_env(monkeypatch: pytest.MonkeyPatch) -> None:
    for var, val in zip(_VARS, ("0", "1", "0", "127.0.0.1", str(_free_port())), strict=True):
        monkeypatch.setenv(var, val)


def test_state_invariant() -> None:
    assert DDPState(rank=0, world_size=2, local_rank=0, is_main=True).is_main
    with pytest.raises(ValueError, match="is_main"):
        DDPState(rank=1, world_size=2, local_rank=1, is_main=True)
    with pytest.raises(ValueError, match="is_main"):
        DDPState(rank=0, world_size=2, local_rank=0, is_main=False)


def test_setup_is_idempotent_and_teardown_is_safe(monkeypatch: pytest.MonkeyPatch) -> None:
    _single_rank_env(monkeypatch)
    cfg = config(ddp={"enabled": True})
    state = setup_ddp(cfg)
    assert (state.rank, state.world_size, state.is_main) == (0, 1, True) and dist.is_initialized()
    assert setup_ddp(cfg) == state
    teardown_ddp()
    assert not dist.is_initialized()
    teardown_ddp()


def test_resolution_prefers_env_then_config(monkeypatch: pytest.MonkeyPatch) -> None:
    with pytest.raises(RuntimeError, match=r"DDP rank not found in env \(RANK\) or config \(ddp.rank\)"):
        setup_ddp(config(ddp={"enabled": True}))
    monkeypatch.setenv("RANK", "zero")
    with pytest.raises(RuntimeError, match="RANK must be an integer"):
        setup_ddp(config(ddp={"enabled": True}))
    monkeypatch.delenv("RANK")
    cfg = config(ddp={"enabled": True, "rank": 0, "world_size": 1, "local_rank": 0,
                      "master_addr": "127.0.0.1", "master_port": _free_port()})
    state = setup_ddp(cfg)  # everything from config; MASTER_* exported for env://
    assert state.world_size == 1 and os.environ["MASTER_ADDR"] == "127.0.0.1"


def test_trainer_takes_ddp_branches_with_a_pretend_second_rank(
    monkeypatch: pytest.MonkeyPatch, tmp_path: Path
) -> None:
    _single_rank_env(monkeypatch)
    cfg = config(tmp_path, ddp={"enabled": True},
                 trainer={"max_steps": 2, "log_every_steps": 1, "eval_every_steps": 2})
    setup_ddp(cfg)
    pretend = DDPState(rank=0, world_size=2, local_rank=0, is_main=True)

    def fake_gather(slots: list[torch.Tensor], mine: torch.Tensor) -> None:
        slots[0].copy_(mine)
        slots[1].copy_(mine + 0.5)

    monkeypatch.setattr(dist, "all_gather", fake_gather)
    monkeypatch.setattr(dist, "all_reduce", lambda t, op=None: t.mul_(2))
    tracker = Mock()
    trainer = Trainer(cfg, run_dir=tmp_path / "r", tracker=tracker, ddp_state=pretend)
    assert isinstance(trainer._model, DistributedDataParallel)
    assert not isinstance(trainer._raw_model, DistributedDataParallel)
    trainer.fit()
    keys = {k for call in tracker.log_metrics.call_args_list for k in call.args[0]}
    assert {"train/loss", "train/loss_rank_0", "train/loss_rank_1", "train/lr_rank_1",
            "train/tokens_total_rank_0", "val/loss", "val/loss_rank_0", "val/loss_rank_1"} <= keys
    r0 = next(c.args[0] for c in tracker.log_metrics.call_args_list if "train/loss_rank_0" in c.args[0])
    r1 = next(c.args[0] for c in tracker.log_metrics.call_args_list if "train/loss_rank_1" in c.args[0])
    assert r1["train/loss_rank_1"] == pytest.approx(r0["train/loss_rank_0"] + 0.5)


def test_non_main_rank_has_no_side_effects(monkeypatch: pytest.MonkeyPatch, tmp_path: Path) -> None:
    _single_rank_env(monkeypatch)
    cfg = config(tmp_path, ddp={"enabled": True}, trainer={"max_steps": 2, "save_every_steps": 1})
    setup_ddp(cfg)
    monkeypatch.setattr(dist, "all_gather", lambda slots, mine: None)
    monkeypatch.setattr(dist, "all_reduce", lambda t, op=None: None)
    tracker = Mock()
    worker = DDPState(rank=1, world_size=2, local_rank=1, is_main=False)
    trainer = Trainer(cfg, run_dir=None, tracker=tracker, ddp_state=worker)
    trainer.fit()
    tracker.log_params.assert_not_called()
    tracker.log_metrics.assert_not_called()
    assert trainer._gather_scalars(x=1.0) is None
    assert trainer._reduce_metrics(a=1.0) == {"a": 1.0}  # the fake all_reduce is an identity


def test_reduce_and_gather_are_identities_without_ddp() -> None:
    trainer = Trainer(config())
    assert trainer._reduce_metrics(a=1.5, b=2.0) == {"a": 1.5, "b": 2.0}
    assert trainer._gather_scalars(a=1.5) == [{"a": 1.5}]


@pytest.mark.slow
def test_two_process_torchrun_gloo_end_to_end(tmp_path: Path) -> None:
    doc = payload(tmp_path / "runs", ddp={"enabled": True},
                  trainer={"max_steps": 6, "micro_batch_size": 2, "grad_accum_steps": 2,
                           "save_every_steps": 3, "log_every_steps": 3, "eval_every_steps": 3})
    path = write_yaml(tmp_path / "ddp.yaml", doc)
    env = dict(os.environ, PYTHONPATH=SRC)
    for var in _VARS:
        env.pop(var, None)
    proc = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
         "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), "-m", "llmtrain",
         "train", "--config", str(path), "--json"],
        capture_output=True, text=True, env=env, timeout=180, check=False)
    assert proc.returncode == 0, proc.stderr[-2000:]
    summary = json.loads(proc.stdout)  # only rank 0 prints
    assert summary["training"]["final_step"] == 6
    runs = list((tmp_path / "runs").iterdir())
    assert len(runs) == 1
    assert len(list((runs[0] / "checkpoints").glob("step_*.pt"))) == 2
