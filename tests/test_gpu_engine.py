"""Native GPT runtime vs the torch-op oracle, and the CUDA trainer path end to end (B200 box)."""

from __future__ import annotations

import copy
import math
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(tmp: Path | None = None, **over):
    from llmtrain.config.schemas import RunConfig

    payload = {
        "run": {"name": "gpu-engine", "device": "cuda"},
        "model": {"name": "gpt", "block_size": 128, "d_model": 128, "n_layers": 2, "n_heads": 2,
                  "d_ff": 512, "dropout": 0.0, "vocab_size": 1000},
        "data": {"name": "synthetic_text", "extra": {"pattern": "arith", "val_examples": 8}},
        "trainer": {"max_steps": 30, "micro_batch_size": 4, "grad_accum_steps": 1,
                    "warmup_steps": 0, "lr": 3e-3, "weight_decay": 0.0, "log_every_steps": 10,
                    "eval_every_steps": 30, "save_every_steps": 10},
        "ddp": {}, "mlflow": {"enabled": False}, "logging": {"log_to_file": False},
        "output": {"root_dir": str(tmp) if tmp else "runs"},
    }
    for section, values in over.items():
        payload[section].update(values)
    return RunConfig.model_validate(payload)


def _model(tie: bool = True, **kw):
    from llmtrain.models.gpt import GPT

    torch.manual_seed(0)
    args = dict(vocab_size=1000, block_size=128, d_model=128, n_layers=2, n_heads=2, d_ff=512,
                dropout=0.0, tie_embeddings=tie)
    args.update(kw)
    return GPT(**args)


@pytest.mark.parametrize("tie", [True, False])
@pytest.mark.parametrize("b,t", [(2, 128), (3, 50)])
def test_native_loss_and_grads_match_torch_ops(tie: bool, b: int, t: int) -> None:
    from llmtrain.models.gpt import GPTAdapter

    oracle = _model(tie).cuda()
    oracle.prefer_native = False
    native = copy.deepcopy(oracle)
    native.prefer_native = True
    adapter = GPTAdapter()
    ids = torch.randint(0, 1000, (b, t), device="cuda")
    labels = torch.randint(0, 1000, (b, t), device="cuda")
    batch = {"input_ids": ids, "labels": labels}

    want, _ = adapter.compute_loss(oracle, batch)
    want.backward()
    got, metrics = adapter.compute_loss(native, batch)
    assert native._engine is not None, "native runtime was not used"
    got.backward()
    assert abs(got.item() - want.item()) < 2e-2 * max(1.0, abs(want.item()))
    assert abs(metrics["loss"] - got.item()) < 1e-6

    ref = dict(oracle.named_parameters())
    for name, p in native.named_parameters():
        assert p.grad is not None, name
        g_ref = ref[name].grad
        denom = g_ref.abs().max().item() + 1e-8
        err = (p.grad - g_ref).abs().max().item() / denom
        assert err < 6e-2, f"{name}: rel err {err:.3e}"


def test_native_logits_and_mask_path() -> None:
    from llmtrain.models.gpt import GPTAdapter

    oracle = _model().cuda().eval()
    oracle.prefer_native = False
    native = copy.deepcopy(oracle)
    native.prefer_native = True
    ids = torch.randint(0, 1000, (2, 100), device="cuda")
    with torch.no_grad():
        want = oracle(ids)
        got = native(ids)
    assert got.shape == want.shape == (2, 100, 1000)
    assert (got - want).abs().max().item() < 5e-2 * want.abs().max().item()

    mask = torch.ones(2, 100, dtype=torch.long, device="cuda")
    mask[0, 70:] = 0
    labels = torch.randint(0, 1000, (2, 100), device="cuda")
    batch = {"input_ids": ids, "labels": labels, "attention_mask": mask}
    adapter = GPTAdapter()
    with torch.no_grad():
        want_loss, _ = adapter.compute_loss(oracle, batch)
        got_loss, _ = adapter.compute_loss(native, batch)
    assert abs(got_loss.item() - want_loss.item()) < 2e-2 * abs(want_loss.item())


def test_state_dict_keys_match_cpu_model_and_alias_is_kept() -> None:
    cpu = _model()
    gpu = copy.deepcopy(cpu).cuda()
    gpu.native_engine()
    assert list(gpu.state_dict().keys()) == list(cpu.state_dict().keys())
    assert gpu.lm_head.weight is gpu.token_embedding.weight
    gpu.load_state_dict(cpu.state_dict())
    assert gpu._engine.bound_to(gpu)
    for (n1, a), (n2, b_) in zip(cpu.state_dict().items(), gpu.state_dict().items(), strict=True):
        assert n1 == n2 and torch.equal(a, b_.cpu()), n1


def test_trainer_native_engine_learns_and_resumes(tmp_path: Path) -> None:
    from llmtrain.engine.b200 import B200TrainEngine
    from llmtrain.training import Trainer

    cfg = _cfg(tmp_path)
    trainer = Trainer(cfg, run_dir=tmp_path / "run")
    assert isinstance(trainer._engine, B200TrainEngine)
    result = trainer.fit()
    assert result.first_step_loss is not None and math.isfinite(result.final_loss)
    assert result.final_loss < result.first_step_loss - 0.5
    assert result.final_val_loss is not None and result.final_val_loss < result.first_step_loss
    assert result.peak_memory > 0
    ckpts = sorted((tmp_path / "run" / "checkpoints").glob("step_*.pt"))
    assert [c.name for c in ckpts] == ["step_000010.pt", "step_000020.pt", "step_000030.pt"]

    payload = torch.load(ckpts[-1], weights_only=False)
    assert all(not v.is_cuda for v in payload["model_state_dict"].values())
    state0 = payload["optimizer_state_dict"]["state"][0]
    assert set(state0) == {"step", "exp_avg", "exp_avg_sq"} and float(state0["step"]) == 30.0

    # resume from step 20 and land (almost) where the uninterrupted run landed
    again = Trainer(cfg, run_dir=tmp_path / "run2")
    resumed = again.fit(resume_from=ckpts[1])
    assert resumed.resumed_from_step == 20
    assert abs(resumed.final_loss - result.final_loss) < 5e-2 * max(1.0, result.final_loss)


def test_checkpoint_from_gpu_loads_into_cpu_model(tmp_path: Path) -> None:
    from llmtrain.training import Trainer

    cfg = _cfg(tmp_path, trainer={"max_steps": 2, "save_every_steps": 2})
    Trainer(cfg, run_dir=tmp_path / "r").fit()
    payload = torch.load(tmp_path / "r" / "checkpoints" / "step_000002.pt", weights_only=False)
    cpu = _model()
    cpu.load_state_dict(payload["model_state_dict"])  # strict: includes causal_mask keys
    opt = torch.optim.AdamW(cpu.parameters())
    opt.load_state_dict(payload["optimizer_state_dict"])


def test_grad_accumulation_matches_big_batch() -> None:
    """accum k over k micro-batches == one batch k times larger (same tokens)."""
    from llmtrain.engine.b200 import GPTEngine

    base = _model().cuda()
    ids = torch.randint(0, 1000, (4, 64), device="cuda")
    labels = torch.randint(0, 1000, (4, 64), device="cuda")

    def grads(chunks: int) -> torch.Tensor:
        m = copy.deepcopy(base)
        eng: GPTEngine = m.native_engine(max_batch=4, max_seq=64)
        g = eng.ensure_grad_arena()
        eng.refresh_shadow()
        acc = torch.zeros(1, device="cuda")
        for part_ids, part_lab in zip(ids.chunk(chunks), labels.chunk(chunks), strict=True):
            scale = 1.0 / (part_ids.numel() * chunks)
            eng.run_forward(part_ids.contiguous(), part_lab.contiguous(), None, mode=2,
                            loss_acc=acc, loss_scale=scale, grad_scale=scale)
            eng.run_backward()
        torch.cuda.synchronize()
        return g.clone()

    g1, g2 = grads(1), grads(2)
    assert (g1 - g2).abs().max().item() < 2e-2 * g1.abs().max().item()


def test_native_matches_oracle_on_gpt2_xl_shaped_layer() -> None:
    """d=1600 / 25 heads / F=6400: none of the GEMM extents is a multiple of the 128x256 tile."""
    from llmtrain.models.gpt import GPTAdapter

    oracle = _model(vocab_size=1000, block_size=256, d_model=1600, n_layers=1, n_heads=25,
                    d_ff=6400).cuda()
    oracle.prefer_native = False
    native = copy.deepcopy(oracle)
    native.prefer_native = True
    adapter = GPTAdapter()
    ids = torch.randint(0, 1000, (2, 200), device="cuda")
    labels = torch.randint(0, 1000, (2, 200), device="cuda")
    batch = {"input_ids": ids, "labels": labels}
    want, _ = adapter.compute_loss(oracle, batch)
    want.backward()
    got, _ = adapter.compute_loss(native, batch)
    got.backward()
    assert abs(got.item() - want.item()) < 2e-2 * abs(want.item())
    ref = dict(oracle.named_parameters())
    for name, p in native.named_parameters():
        g_ref = ref[name].grad
        err = (p.grad - g_ref).abs().max().item() / (g_ref.abs().max().item() + 1e-8)
        assert err < 8e-2, f"{name}: rel err {err:.3e}"


@pytest.mark.parametrize("d_model,n_heads", [(64, 4), (256, 8), (96, 2)])
def test_narrow_heads_run_natively_and_match_torch_ops(d_model: int, n_heads: int) -> None:
    """head_dim 16 / 32 / 48 (every preset of the reference uses 16 or 32): the engine pads the
    heads to 64 lanes; losses and ALL parameter gradients must still match the torch-op oracle."""
    from llmtrain.engine.b200 import supports
    from llmtrain.models.gpt import GPTAdapter

    oracle = _model(True, d_model=d_model, n_heads=n_heads, d_ff=4 * d_model, block_size=64).cuda()
    assert supports(oracle)
    oracle.prefer_native = False
    native = copy.deepcopy(oracle)
    native.prefer_native = True
    adapter = GPTAdapter()
    ids = torch.randint(0, 1000, (3, 64), device="cuda")
    labels = torch.randint(0, 1000, (3, 64), device="cuda")
    mask = torch.ones(3, 64, dtype=torch.long, device="cuda")
    mask[1, 40:] = 0
    batch = {"input_ids": ids, "labels": labels, "attention_mask": mask}

    want, _ = adapter.compute_loss(oracle, batch)
    want.backward()
    got, _ = adapter.compute_loss(native, batch)
    assert native._engine is not None, "native runtime was not used"
    got.backward()
    assert abs(got.item() - want.item()) < 2e-2 * max(1.0, abs(want.item()))
    ref = dict(oracle.named_parameters())
    for name, p in native.named_parameters():
        g_ref = ref[name].grad
        err = (p.grad - g_ref).abs().max().item() / (g_ref.abs().max().item() + 1e-8)
        assert err < 6e-2, f"{name}: rel err {err:.3e}"
    # two accumulated backward passes double the gradients (the padded temporaries are re-zeroed)
    got2, _ = adapter.compute_loss(native, batch)
    got2.backward()
    w = native.blocks[0].attn.qkv_proj.weight
    err = (w.grad - 2 * ref["blocks.0.attn.qkv_proj.weight"].grad).abs().max().item()
    assert err / (ref["blocks.0.attn.qkv_proj.weight"].grad.abs().max().item() + 1e-8) < 0.12


def test_trainer_runs_on_device_resident_synthetic_tokens(tmp_path: Path) -> None:
    """``data.extra.device_resident``: batches come from the library's token kernel,
    nothing is copied from the host; the trainer contract (TrainResult, finite loss near ln V for
    random tokens) is unchanged."""
    from llmtrain.engine.b200 import B200TrainEngine
    from llmtrain.training import Trainer

    cfg = _cfg(tmp_path,
               data={"extra": {"pattern": "random", "device_resident": True}},
               trainer={"max_steps": 5, "eval_every_steps": 1000, "save_every_steps": 1000})
    trainer = Trainer(cfg)
    assert isinstance(trainer._engine, B200TrainEngine)
    first = next(iter(trainer._train_loader))
    assert first["input_ids"].is_cuda and first["input_ids"].shape == (4, 128)
    assert torch.equal(first["labels"][:, :-1], first["input_ids"][:, 1:])
    result = trainer.fit()
    assert result.final_step == 5
    assert abs(result.final_loss - math.log(1000)) < 1.0


def test_deterministic_kernels_make_resume_reproduce_the_uninterrupted_run(tmp_path: Path) -> None:
    """``trainer.extra.deterministic_kernels`` + dropout: a run resumed from step 4 ends on the loss
    of the uninterrupted run to the reference's own tolerance (its CPU test demands 1e-5,
    `tests/test_checkpoint.py:301-320`), and two identical runs agree bit for bit."""
    from llmtrain.ops import native
    from llmtrain.training import Trainer

    over = dict(model={"dropout": 0.1},
                trainer={"max_steps": 8, "save_every_steps": 4, "eval_every_steps": 1000,
                         "grad_accum_steps": 2, "extra": {"deterministic_kernels": True}})
    try:
        torch.manual_seed(7)
        first = Trainer(_cfg(tmp_path, **over), run_dir=tmp_path / "a")
        whole = first.fit()
        weights = first._engine._gpt.p32.clone()

        torch.manual_seed(7)
        second = Trainer(_cfg(tmp_path, **over), run_dir=tmp_path / "b")
        again = second.fit()
        assert again.final_loss == whole.final_loss
        assert torch.equal(second._engine._gpt.p32, weights)

        torch.manual_seed(99)  # a different process state: everything must come from the checkpoint
        third = Trainer(_cfg(tmp_path, **over), run_dir=tmp_path / "c")
        resumed = third.fit(resume_from=tmp_path / "a" / "checkpoints" / "step_000004.pt")
        assert resumed.resumed_from_step == 4
        assert abs(resumed.final_loss - whole.final_loss) <= 1e-5
        assert torch.equal(third._engine._gpt.p32, weights)
    finally:
        native.lib().llmtrain_b200_set_deterministic(0)


def test_padding_mask_semantics_match_the_reference_at_every_position() -> None:
    """With an ``attention_mask`` the attention branch of padded QUERY rows is dropped before the
    residual add (reference `src/llmtrain/models/gpt.py:73-74`): logits agree with the torch-op
    oracle at padded positions too, and so do the parameter gradients of a masked loss."""
    from llmtrain.models.gpt import GPTAdapter

    oracle = _model().cuda()
    oracle.prefer_native = False
    native = copy.deepcopy(oracle)
    native.prefer_native = True
    ids = torch.randint(0, 1000, (2, 96), device="cuda")
    mask = torch.ones(2, 96, dtype=torch.long, device="cuda")
    mask[0, 60:] = 0
    mask[1, 90:] = 0
    oracle.eval()
    native.eval()
    with torch.no_grad():
        want = oracle(ids, attention_mask=mask)
        got = native(ids, attention_mask=mask)
    assert (got - want).abs().max().item() < 5e-2 * want.abs().max().item()

    oracle.train()
    native.train()
    labels = torch.randint(0, 1000, (2, 96), device="cuda")
    batch = {"input_ids": ids, "labels": labels, "attention_mask": mask}
    adapter = GPTAdapter()
    want_loss, _ = adapter.compute_loss(oracle, batch)
    want_loss.backward()
    got_loss, _ = adapter.compute_loss(native, batch)
    got_loss.backward()
    assert abs(got_loss.item() - want_loss.item()) < 2e-2 * abs(want_loss.item())
    for (name, p), (_, q) in zip(native.named_parameters(), oracle.named_parameters(), strict=True):
        denom = q.grad.abs().max().item() + 1e-8
        err = (p.grad - q.grad).abs().max().item() / denom
        assert err < 6e-2, f"{name}: rel err {err:.3e}"
