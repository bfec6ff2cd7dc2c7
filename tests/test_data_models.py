"""Data modules and the CPU (oracle) model path."""

from __future__ import annotations

import copy
from pathlib import Path

import pytest
import torch

from llmtrain.models.gpt import GPT, CausalSelfAttention, GPTAdapter
from tests.helpers import TINY_GPT, config


def _gpt(**kw) -> GPT:
    args = dict(vocab_size=32, block_size=16, d_model=64, n_layers=2, n_heads=4, d_ff=128,
                dropout=0.0)
    args.update(kw)
    return GPT(**args)


# ----------------------------------------------------------------------------- data
def test_dummy_text_shapes_determinism_and_limits() -> None:
    from llmtrain.data.dummy_text import DummyTextDataModule

    dm = DummyTextDataModule()
    with pytest.raises(RuntimeError, match="setup must be called"):
        dm.train_dataloader()
    cfg = config(trainer={"max_steps": 500, "micro_batch_size": 2}, model={"vocab_size": 50})
    dm.setup(cfg)
    loader = dm.train_dataloader()
    batch = next(iter(loader))
    assert batch["input_ids"].shape == (2, 8) and batch["input_ids"].dtype == torch.long
    assert torch.equal(batch["labels"], batch["input_ids"])  # deliberately unshifted
    assert int(batch["attention_mask"].min()) == 1
    assert len(loader.dataset) == 128 and len(dm.val_dataloader().dataset) == 25
    again = next(iter(dm.train_dataloader()))
    assert torch.equal(again["input_ids"], batch["input_ids"])


def test_synthetic_text_full_length_shifted_rows() -> None:
    from llmtrain.data.synthetic_text import SyntheticTextDataModule

    cfg = config(model={**TINY_GPT, "block_size": 64, "vocab_size": 100},
                 data={"name": "synthetic_text", "extra": {"num_examples": 10, "val_examples": 4}},
                 trainer={"micro_batch_size": 3})
    dm = SyntheticTextDataModule()
    dm.setup(cfg)
    batch = next(iter(dm.train_dataloader()))
    assert batch["input_ids"].shape == (3, 64)
    assert torch.equal(batch["labels"][:, :-1], batch["input_ids"][:, 1:])
    assert int(batch["input_ids"].max()) < 100
    assert len(dm.val_dataloader().dataset) == 4

    arith = config(model={**TINY_GPT, "block_size": 16}, trainer={"micro_batch_size": 1},
                   data={"name": "synthetic_text", "extra": {"pattern": "arith"}})
    dm.setup(arith)
    row = next(iter(dm.train_dataloader()))
    step = (row["labels"][0] - row["input_ids"][0]) % 16
    assert len(step.unique()) == 1 and dm.val_dataloader() is None
    with pytest.raises(ValueError, match="pattern"):
        dm.setup(config(model=TINY_GPT, data={"name": "synthetic_text", "extra": {"pattern": "x"}}))


class _ToyTokenizer:
    def encode(self, text: str) -> list[int]:
        return [ord(c) % 50 for c in text]


def test_hf_text_pipeline_cache_and_sampler_hints(tmp_path: Path, monkeypatch) -> None:
    datasets = pytest.importorskip("datasets")
    from torch.utils.data.distributed import DistributedSampler

    from llmtrain.data.hf_text import HFTextDataModule

    calls = {"n": 0}

    def fake_load_dataset(name, variant, split, cache_dir):  # noqa: ANN001, ANN202
        calls["n"] += 1
        return datasets.Dataset.from_dict({"text": ["abcdefghij" * 5, None, "klmnopqrstuvwxyz" * 4]})

    monkeypatch.setattr("datasets.load_dataset", fake_load_dataset)
    cfg = config(
        model={"block_size": 8},
        data={"name": "hf_text", "dataset_name": "toy/set", "dataset_config": "v1",
              "text_column": "text", "cache_dir": str(tmp_path / "cache"), "num_workers": 0},
        trainer={"micro_batch_size": 2},
    )
    dm = HFTextDataModule()
    with pytest.raises(ValueError, match="tokenizer"):
        dm.setup(cfg, tokenizer=None)
    dm.setup(cfg, tokenizer=_ToyTokenizer())
    assert calls["n"] == 2  # train + validation
    batch = next(iter(dm.train_dataloader()))
    assert batch["input_ids"].shape == (2, 8) and batch["labels"].dtype == torch.long
    assert torch.equal(batch["labels"][:, :-1], batch["input_ids"][:, 1:])  # one-token shift
    assert (tmp_path / "cache" / "processed" / "toy__set__v1__b8__train").exists()

    monkeypatch.setattr("datasets.load_dataset", lambda *a, **k: pytest.fail("cache not used"))
    HFTextDataModule().setup(cfg, tokenizer=_ToyTokenizer())

    hinted = cfg.model_copy(update={"ddp": cfg.ddp.model_copy(update={"world_size": 4, "rank": 2})})
    monkeypatch.setattr("llmtrain.data.hf_text.dist.is_initialized", lambda: False)
    dm2 = HFTextDataModule()
    dm2.setup(hinted, tokenizer=_ToyTokenizer())
    sampler = dm2.train_dataloader().sampler
    assert isinstance(sampler, DistributedSampler)
    assert sampler.num_replicas == 4 and sampler.rank == 2


# ----------------------------------------------------------------------------- model
def test_attention_and_model_are_causal() -> None:
    torch.manual_seed(0)
    attn = CausalSelfAttention(64, 4, 16, 0.0).eval()
    x = torch.randn(2, 10, 64)
    y = x.clone()
    y[:, 6:] += 1.0
    assert torch.allclose(attn(x)[:, :6], attn(y)[:, :6], atol=1e-6)

    model = _gpt().eval()
    ids = torch.randint(0, 32, (2, 12))
    other = ids.clone()
    other[:, 8:] = (other[:, 8:] + 1) % 32
    assert torch.allclose(model(ids)[:, :8], model(other)[:, :8], atol=1e-6)
    assert model(ids).shape == (2, 12, 32)
    with pytest.raises(ValueError, match="exceeds block size"):
        model(torch.zeros(1, 17, dtype=torch.long))


def test_padding_mask_semantics_fully_masked_rows_stay_finite() -> None:
    attn = CausalSelfAttention(64, 4, 16, 0.0).eval()
    x = torch.randn(1, 6, 64)
    mask = torch.tensor([[0, 0, 1, 1, 1, 0]])
    out = attn(x, attention_mask=mask)
    assert torch.isfinite(out).all()
    assert out[0, 0].abs().max() == 0 and out[0, 5].abs().max() == 0  # padded queries zeroed
    with pytest.raises(ValueError, match="attention_mask"):
        attn(x, attention_mask=torch.ones(1, 5))


def test_weight_tying_init_and_parameter_count() -> None:
    tied, untied = _gpt(), _gpt(tie_embeddings=False)
    assert tied.lm_head.weight is tied.token_embedding.weight
    assert untied.lm_head.weight is not untied.token_embedding.weight
    n = sum(p.numel() for p in tied.parameters())
    d, f, v, t, layers = 64, 128, 32, 16, 2
    per_layer = 4 * d + (3 * d * d + 3 * d) + (d * d + d) + (f * d + f) + (d * f + d)
    assert n == v * d + t * d + layers * per_layer + 2 * d
    assert len(list(tied.parameters())) == 2 + 12 * layers + 2
    big = _gpt(n_layers=8)
    assert big.blocks[0].mlp_proj.weight.std() < 0.6 * big.blocks[0].mlp_fc.weight.std()


def test_state_dict_layout_matches_reference_keys_and_round_trips() -> None:
    model = _gpt()
    keys = list(model.state_dict())
    assert keys[:2] == ["token_embedding.weight", "position_embedding.weight"]
    assert keys[2:15] == [
        "blocks.0.ln_1.weight", "blocks.0.ln_1.bias", "blocks.0.attn.causal_mask",
        "blocks.0.attn.qkv_proj.weight", "blocks.0.attn.qkv_proj.bias",
        "blocks.0.attn.out_proj.weight", "blocks.0.attn.out_proj.bias", "blocks.0.ln_2.weight",
        "blocks.0.ln_2.bias", "blocks.0.mlp_fc.weight", "blocks.0.mlp_fc.bias",
        "blocks.0.mlp_proj.weight", "blocks.0.mlp_proj.bias"]
    assert keys[-3:] == ["ln_f.weight", "ln_f.bias", "lm_head.weight"]
    mask = model.state_dict()["blocks.1.attn.causal_mask"]
    assert mask.shape == (1, 1, 16, 16) and mask.dtype == torch.bool
    assert bool(mask[0, 0, 0, 1]) and not bool(mask[0, 0, 1, 0])
    assert "blocks.0.attn.causal_mask" not in dict(model.named_buffers())  # never a live buffer

    clone = _gpt()
    clone.load_state_dict(copy.deepcopy(model.state_dict()))  # strict
    ids = torch.randint(0, 32, (1, 8))
    assert torch.equal(clone.eval()(ids), model.eval()(ids))
    sd = model.state_dict()
    sd.pop("blocks.0.attn.causal_mask")
    clone.load_state_dict(sd)  # files without the mask keys load too
    with pytest.raises(RuntimeError, match="size mismatch"):
        _gpt(block_size=8).load_state_dict(model.state_dict())


def test_adapter_loss_paths_and_validation() -> None:
    adapter = GPTAdapter()
    cfg = config(model=TINY_GPT)
    model = adapter.build_model(cfg)
    assert isinstance(model, GPT) and model.vocab_size == 16
    ids = torch.randint(0, 16, (3, 8))
    loss, metrics = adapter.compute_loss(model, {"input_ids": ids, "labels": ids})
    assert loss.requires_grad and abs(metrics["loss"] - loss.item()) < 1e-6
    mask = torch.ones_like(ids)
    mask[:, 5:] = 0
    masked, _ = adapter.compute_loss(model, {"input_ids": ids, "labels": ids, "attention_mask": mask})
    assert masked.item() != pytest.approx(loss.item())
    with pytest.raises(ValueError, match="no valid target tokens"):
        adapter.compute_loss(model, {"input_ids": ids, "labels": ids,
                                     "attention_mask": torch.zeros_like(ids)})
    with pytest.raises(ValueError, match="torch.long"):
        adapter.compute_loss(model, {"input_ids": ids.int(), "labels": ids})
    with pytest.raises(ValueError, match="same shape"):
        adapter.compute_loss(model, {"input_ids": ids, "labels": ids[:, :4]})
    with pytest.raises(ValueError, match=">= 2"):
        adapter.compute_loss(model, {"input_ids": ids[:, :1], "labels": ids[:, :1]})


def test_tokenizer_is_offline_safe() -> None:
    tok = GPTAdapter().build_tokenizer(config(model=TINY_GPT))
    assert tok.n_vocab == 50257
    model = GPTAdapter().build_model(config(model={**TINY_GPT, "vocab_size": None}))
    assert model.token_embedding.num_embeddings == 50257


def test_dummy_gpt_is_small_and_ignores_mask_in_loss() -> None:
    from llmtrain.models.dummy_gpt import DummyGPTAdapter

    adapter = DummyGPTAdapter()
    model = adapter.build_model(config(model={"d_model": 384, "n_heads": 6}))
    assert model.embed.embedding_dim == 64 and model.embed.num_embeddings == 128
    assert adapter.build_tokenizer(config()) is None
    ids = torch.randint(0, 128, (2, 8))
    model.eval()
    a, _ = adapter.compute_loss(model, {"input_ids": ids, "labels": ids})
    assert torch.isfinite(a)


def test_generation_greedy_sampling_and_checkpoint_loader(tmp_path: Path) -> None:
    from llmtrain.generation import generate, load_model_from_checkpoint
    from llmtrain.training import Trainer

    cfg = config(tmp_path, model=TINY_GPT,
                 trainer={"max_steps": 60, "micro_batch_size": 4, "lr": 3e-3, "weight_decay": 0.0,
                          "save_every_steps": 60, "log_every_steps": 60, "eval_every_steps": 60})
    Trainer(cfg, run_dir=tmp_path / "g").fit()
    model, loaded_cfg = load_model_from_checkpoint(tmp_path / "g" / "checkpoints" / "step_000060.pt")
    assert loaded_cfg.model.vocab_size == 16 and not model.training
    prompt = torch.tensor([[3, 7]])
    greedy = generate(model, prompt, max_new_tokens=5, temperature=0.0)
    assert greedy.shape == (1, 7) and torch.equal(greedy[:, :2], prompt)
    # the copy task teaches "repeat the current token": greedy continuation repeats the last one
    assert greedy[0, 2:].tolist() == [7] * 5
    g1 = generate(model, prompt, max_new_tokens=12, temperature=1.0, top_k=3,
                  generator=torch.Generator().manual_seed(0))
    g2 = generate(model, prompt, max_new_tokens=12, temperature=1.0, top_k=3,
                  generator=torch.Generator().manual_seed(0))
    assert torch.equal(g1, g2) and g1.shape == (1, 14) and int(g1.max()) < 16
    with pytest.raises(ValueError, match="batch, time"):
        generate(model, torch.tensor([1, 2]), max_new_tokens=1)
