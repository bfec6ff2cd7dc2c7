"""Reference arm of bench.py: drives the UNMODIFIED reference package (``baseline/_ref``).

Nothing of this repo's ``llmtrain`` (models, kernels, engine) is imported here — ``bench.py`` puts
``baseline/_ref`` first on ``sys.path`` and does not add ``src/``.  The reference is exercised
through its own public API and stock code path:

* its ``Trainer`` / ``Trainer.fit`` (`src/llmtrain/training/trainer.py:56-548`), stock ``GPT`` model
  and ``GPTAdapter.compute_loss``, stock ``torch.optim.AdamW`` + ``clip_grad_norm_`` +
  ``DistributedDataParallel``;
* its plugin registry (`src/llmtrain/registry/data.py`) to add a ``synthetic_text`` data module:
  the shipped ``dummy_text`` cannot produce sequences longer than 8 tokens and ``hf_text``
  downloads, so a registered plugin is the reference's own extension point for seq-1024 data.

Three things the reference cannot do by itself are supplied from the *outside*, without touching
its source (all recorded in the emitted JSON under ``notes.reference_mode``):

1. ``run.device`` / ``ddp.backend`` are ``Literal["cpu","mps"]`` / ``Literal["gloo"]`` in its
   schema; the config object is assembled with pydantic's ``model_construct`` so that the stock
   trainer places the model on ``cuda`` and DDP runs over ``cpu:gloo,cuda:nccl``.
2. bf16: ``Trainer.fit`` is called inside ``torch.autocast("cuda", dtype=torch.bfloat16)`` (the
   reference has no mixed-precision switch; fp32 master weights and AdamW stay as it has them).
3. ``tiktoken.get_encoding`` needs the network; it is replaced by a stub exposing ``n_vocab`` (the
   benchmark never tokenises).
"""

from __future__ import annotations

from typing import Any


def _install_plugins(seq_len: int, vocab: int, examples: int) -> None:
    import tiktoken
    import torch
    from llmtrain.data.base import DataModule
    from llmtrain.registry.data import available_data_modules, register_data_module
    from torch.utils.data import DataLoader, Dataset

    class _StubEncoding:
        n_vocab = 50257

        def encode(self, text: str) -> list[int]:
            raise RuntimeError("offline benchmark: tokenizer is never used")

    tiktoken.get_encoding = lambda name: _StubEncoding()  # type: ignore[assignment]

    class _Rows(Dataset):
        def __len__(self) -> int:
            return examples

        def __getitem__(self, index: int) -> dict[str, torch.Tensor]:
            gen = torch.Generator().manual_seed(1234 * 1_000_003 + index)
            stream = torch.randint(0, vocab, (seq_len + 1,), dtype=torch.long, generator=gen)
            ids = stream[:-1].clone()
            return {"input_ids": ids, "labels": stream[1:].clone(),
                    "attention_mask": torch.ones_like(ids)}

    if "synthetic_text" in available_data_modules():
        return

    @register_data_module("synthetic_text")
    class SyntheticText(DataModule):
        def setup(self, cfg: Any, tokenizer: Any | None = None) -> None:
            self._mb = cfg.trainer.micro_batch_size

        def train_dataloader(self) -> DataLoader:
            return DataLoader(_Rows(), batch_size=self._mb, num_workers=0, shuffle=False,
                              pin_memory=True)

        def val_dataloader(self) -> DataLoader | None:
            return None


def _config(model: dict[str, Any], mb: int, accum: int, world: int, steps: int) -> Any:
    from llmtrain.config import schemas as S

    return S.RunConfig.model_construct(
        schema_version=1,
        run=S.RunSectionConfig.model_construct(name="bench-ref", seed=1234, device="cuda",
                                               deterministic=True, notes=None),
        model=S.ModelConfig(name="gpt", **model),
        data=S.DataConfig(name="synthetic_text", num_workers=0),
        trainer=S.TrainerConfig(max_steps=10**6, micro_batch_size=mb, grad_accum_steps=accum,
                                lr=3e-4, weight_decay=0.1, warmup_steps=0, max_grad_norm=1.0,
                                log_every_steps=1, eval_every_steps=10**9,
                                save_every_steps=10**9),
        ddp=S.DDPConfig.model_construct(enabled=world > 1, backend="cpu:gloo,cuda:nccl",
                                        init_method="env://", timeout_sec=1800,
                                        find_unused_parameters=False, rank=None, world_size=None,
                                        local_rank=None, master_addr=None, master_port=None),
        mlflow=S.MLflowConfig(enabled=False),
        logging=S.LoggingConfig(level="WARNING", log_to_file=False),
        output=S.OutputConfig(),
    )


def run(args: Any, *, rank: int, local_rank: int, world: int, model: dict[str, Any], metric: str,
        timed: Any, max_over_ranks: Any, sampler: Any, config: dict[str, Any]) -> dict[str, Any]:
    import torch

    import llmtrain
    from llmtrain.distributed import setup_ddp, teardown_ddp
    from llmtrain.training import Trainer

    assert "baseline/_ref" in llmtrain.__file__.replace("\\", "/"), llmtrain.__file__
    torch.cuda.set_device(local_rank % torch.cuda.device_count())
    seq, mb, accum = args.seq_len, args.micro_batch, args.grad_accum
    _install_plugins(seq, model["vocab_size"], max(256, 2 * mb * accum))
    cfg = _config(model, mb, accum, world, args.steps + args.warmup)
    ddp_state = setup_ddp(cfg) if world > 1 else None
    try:
        trainer = Trainer(cfg, ddp_state=ddp_state)
        tokens_per_step = mb * accum * seq * world

        def fit(n: int) -> None:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                trainer.fit(max_steps_override=n)

        fit(args.warmup)
        with sampler(torch.cuda.current_device()) as clocks:
            ms = timed(lambda: fit(args.steps), world)
        ms = max_over_ranks(ms, world)
        value = tokens_per_step * args.steps / (ms / 1e3)
        h2d = accum * (3 * mb * seq * 8)  # ids + labels + mask, int64, moved by _move_batch
        out = {
            "metric": metric,
            "value": value,
            "unit": "tokens/s",
            "n_gpus": world,
            "steps": args.steps,
            "warmup": args.warmup,
            "ms_per_step": ms / args.steps,
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None,
            "dtype": "bf16",
            "data": "synthetic",
            "impl": "reference",
            "config": config,
            "notes": {"reference_mode": "unmodified baseline/_ref: stock Trainer.fit/GPT/AdamW/DDP; "
                                        "device=cuda + backend=cpu:gloo,cuda:nccl via pydantic "
                                        "model_construct; outer torch.autocast(bf16); "
                                        "synthetic_text registered through the reference's plugin "
                                        "registry; tiktoken stubbed (offline)"},
            "clocks": clocks.summary(),
            "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4 * accum,
                    "api": "reference llmtrain.training.Trainer.fit (its only path is end-to-end: "
                           "blocking H2D per micro-batch, loss.item() per micro-batch)"},
            "gpu_launches": 0,
            "peak_memory_gb": torch.cuda.max_memory_allocated() / 2**30,
        }
        return out if rank == 0 else {}
    finally:
        if ddp_state is not None:
            teardown_ddp()
