"""B200-native execution: the C++ GPT runtime bound to a :class:`llmtrain.models.gpt.GPT`.

Two layers live here:

* :class:`GPTEngine` — owns the flat arenas (fp32 master weights that the module's
  ``nn.Parameter`` objects alias, fp32 gradient arena, bf16 weight shadow, activation arena) and
  drives ``llmtrain_b200_gpt_*`` (``csrc/engine/gpt_engine.cu``).  It backs ``GPT.forward`` and
  ``GPTAdapter.compute_loss`` on CUDA so the public model API keeps working, with gradients
  delivered through a single whole-model ``autograd.Function``.
* :class:`B200TrainEngine` — the :class:`~llmtrain.engine.base.TrainEngine` the trainer uses on
  CUDA.  It replaces the reference's autograd + ``DistributedDataParallel`` + ``clip_grad_norm_``
  + ``AdamW.step`` sequence (`src/llmtrain/training/trainer.py:362-395`) with: fused
  forward/backward per micro-batch (loss stays on the device), gradient reduction across ranks
  (NVLink peer-memory kernels, :mod:`llmtrain.parallel.nvl`; NCCL as the cross-node fallback),
  and one fused clip + AdamW + shadow-refresh + grad-zero pass.  A real ``torch.optim.AdamW``
  is kept purely as the *state container* so ``optimizer.state_dict()`` is produced by torch and
  checkpoints stay byte-compatible with the reference.
"""

from __future__ import annotations

import ctypes
import logging
import os
from collections.abc import Callable
from typing import Any

import torch
from torch import nn

from llmtrain.config.schemas import RunConfig
from llmtrain.distributed import DDPState
from llmtrain.engine.base import Batch, TrainEngine
from llmtrain.ops import kernels as K
from llmtrain.ops import native
from llmtrain.ops.native import check, ptr, stream_handle

logger = logging.getLogger(__name__)

_LAYER_FIELDS = (
    ("ln_1", "weight"), ("ln_1", "bias"),
    ("attn.qkv_proj", "weight"), ("attn.qkv_proj", "bias"),
    ("attn.out_proj", "weight"), ("attn.out_proj", "bias"),
    ("ln_2", "weight"), ("ln_2", "bias"),
    ("mlp_fc", "weight"), ("mlp_fc", "bias"),
    ("mlp_proj", "weight"), ("mlp_proj", "bias"),
)


def supports(model: nn.Module) -> bool:
    """Can the native runtime execute this module?

    GPT on CUDA with heads of width 8, 16, ..., 64: the attention kernels are written for 64-wide
    heads and the engine runs narrower ones zero-padded (``csrc/kernels/headpad.cu``), which covers
    every preset of the reference (16- and 32-wide heads) as well as the GPT-2 family.
    """
    from llmtrain.models.gpt import GPT

    return (
        isinstance(model, GPT)
        and model.d_model % model.n_heads == 0
        and model.d_model // model.n_heads <= 64
        and (model.d_model // model.n_heads) % 8 == 0
        and model.d_model % 8 == 0
        and model.d_ff % 8 == 0
        and model.d_model <= 2048
    )


def _typed_lib() -> ctypes.CDLL:
    lib = native.lib()
    if not getattr(lib, "_gpt_typed", False):
        for name in (
            "llmtrain_b200_gpt_param_elems",
            "llmtrain_b200_gpt_shadow_elems",
            "llmtrain_b200_gpt_activation_bytes",
            "llmtrain_b200_gpt_logits_offset",
            "llmtrain_b200_gpt_loss_rows_offset",
        ):
            fn = getattr(lib, name)
            fn.restype = ctypes.c_size_t
            fn.argtypes = [ctypes.c_void_p]
        lib.llmtrain_b200_gpt_param_offset.restype = ctypes.c_longlong
        lib.llmtrain_b200_gpt_param_offset.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.llmtrain_b200_gpt_padded_vocab.argtypes = [ctypes.c_void_p]
        lib.llmtrain_b200_gpt_destroy.argtypes = [ctypes.c_void_p]
        lib.llmtrain_b200_gpt_destroy.restype = None
        lib._gpt_typed = True
    return lib


def _sub(module: nn.Module, dotted: str) -> nn.Module:
    for part in dotted.split("."):
        module = getattr(module, part)
    return module



_TRACE_FIT = os.environ.get("LLMTRAIN_FIT_TRACE") is not None  # Trainer.fit host/device timeline


class StepScalars:
    """Loss and gradient norm of one optimiser step on their way to the host.

    ``train_step`` enqueues a device-to-pinned-host copy of both and an event behind it; reading
    (``float(x)``, ``x.grad_norm``) waits for THAT event only, so a caller may enqueue the next
    step first and read this one while the GPU is busy (``Trainer.fit`` does).
    """

    __slots__ = ("_row", "_event", "_value", "timing_event")

    def __init__(self, row: torch.Tensor, event: torch.cuda.Event) -> None:
        self._row = row
        self._event: torch.cuda.Event | None = event
        self._value: tuple[float, float] | None = None

    def _read(self) -> tuple[float, float]:
        if self._value is None:
            assert self._event is not None
            self._event.synchronize()
            self._value = (float(self._row[0]), float(self._row[1]))
            self._event = None  # the ring slot may be reused from here on
        return self._value

    def __float__(self) -> float:
        return self._read()[0]

    def item(self) -> float:
        return self._read()[0]

    @property
    def grad_norm(self) -> float:
        return self._read()[1]


class GPTEngine:
    """Memory plan + kernel sequencing for one GPT module on one B200."""

    def __init__(self, model: nn.Module, *, max_batch: int = 1, max_seq: int | None = None) -> None:
        if not supports(model):
            raise RuntimeError(
                "The sm_100a runtime needs a GPT with head_dim (d_model / n_heads) in 8, 16, ..., 64, "
                "d_model <= 2048 and dims divisible by 8; "
                f"got d_model={getattr(model, 'd_model', '?')} n_heads={getattr(model, 'n_heads', '?')}. "
                "Use trainer.extra.engine: torch for other shapes."
            )
        self._model_id = id(model)
        self.model = model
        self.device = model.token_embedding.weight.device
        self._lib = _typed_lib()
        self._handle = ctypes.c_void_p()
        self._plan = (0, 0)
        self._act: torch.Tensor | None = None
        self.p32: torch.Tensor | None = None
        self.g32: torch.Tensor | None = None
        self.p16: torch.Tensor | None = None
        self.shadow_fresh = False
        self._live: tuple[Any, ...] = ()
        self._dropout_seed: int | None = None
        self._dropout_calls = 0
        self._dropout_armed = False
        self._loss_acc = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._replan(max_batch, max_seq or model.block_size)
        self._adopt_parameters()

    # ------------------------------------------------------------------ lifetime
    def bound_to(self, model: nn.Module) -> bool:
        if id(model) != self._model_id or self.p32 is None:
            return False
        w = model.token_embedding.weight
        lo, hi = self.p32.data_ptr(), self.p32.data_ptr() + self.p32.numel() * 4
        return w.is_cuda and lo <= w.data_ptr() < hi

    def __del__(self) -> None:  # pragma: no cover - best effort
        try:
            if self._handle:
                self._lib.llmtrain_b200_gpt_destroy(self._handle)
        except Exception:
            pass

    def _replan(self, max_batch: int, max_seq: int) -> None:
        m = self.model
        max_seq = min(max_seq, m.block_size)
        if self._handle:
            self._lib.llmtrain_b200_gpt_destroy(self._handle)
            self._handle = ctypes.c_void_p()
        handle = ctypes.c_void_p()
        check(
            self._lib.llmtrain_b200_gpt_create(
                m.vocab_size, m.block_size, m.d_model, m.n_layers, m.n_heads, m.d_ff,
                1 if m.tie_embeddings else 0, max_batch, max_seq, ctypes.byref(handle),
            ),
            "gpt_create",
        )
        self._handle = handle
        self._dropout_armed = False  # a fresh runtime handle starts with dropout off
        self._plan = (max_batch, max_seq)
        self.n_params = int(self._lib.llmtrain_b200_gpt_param_elems(handle))
        self.n_shadow = int(self._lib.llmtrain_b200_gpt_shadow_elems(handle))
        self.padded_vocab = int(self._lib.llmtrain_b200_gpt_padded_vocab(handle))
        act_bytes = int(self._lib.llmtrain_b200_gpt_activation_bytes(handle))
        self._act = None  # release before allocating the bigger arena
        self._act = torch.empty(act_bytes, dtype=torch.uint8, device=self.device)
        self._logits_off = int(self._lib.llmtrain_b200_gpt_logits_offset(handle))
        self._loss_rows_off = int(self._lib.llmtrain_b200_gpt_loss_rows_offset(handle))
        if self.p32 is not None:
            self._bind()

    def _bind(self) -> None:
        assert self.p32 is not None and self.p16 is not None and self._act is not None
        g32 = self.g32 if self.g32 is not None else self.p32  # forward-only use never touches g32
        check(
            self._lib.llmtrain_b200_gpt_bind(
                self._handle, ptr(self.p32), ptr(g32), ptr(self.p16), ptr(self._act),
                ctypes.c_size_t(self._act.numel()),
            ),
            "gpt_bind",
        )

    def _offset(self, layer: int, which: int) -> int:
        off = int(self._lib.llmtrain_b200_gpt_param_offset(self._handle, layer, which))
        if off < 0:
            raise RuntimeError(f"no arena offset for layer={layer} which={which}")
        return off

    def named_slots(self) -> list[tuple[str, nn.Parameter, int]]:
        """``(qualified name, parameter, arena offset)`` for every distinct parameter."""
        m = self.model
        slots: list[tuple[str, nn.Parameter, int]] = [
            ("position_embedding.weight", m.position_embedding.weight, self._offset(-1, 0)),
            ("ln_f.weight", m.ln_f.weight, self._offset(-1, 1)),
            ("ln_f.bias", m.ln_f.bias, self._offset(-1, 2)),
            ("token_embedding.weight", m.token_embedding.weight, self._offset(-1, 3)),
        ]
        if not m.tie_embeddings:
            slots.append(("lm_head.weight", m.lm_head.weight, self._offset(-1, 4)))
        for i, block in enumerate(m.blocks):
            for which, (sub, attr) in enumerate(_LAYER_FIELDS):
                param = getattr(_sub(block, sub), attr)
                slots.append((f"blocks.{i}.{sub}.{attr}", param, self._offset(i, which)))
        return slots

    def _adopt_parameters(self) -> None:
        """Move every parameter into the flat fp32 arena; the module's tensors become views."""
        self.p32 = torch.zeros(self.n_params, dtype=torch.float32, device=self.device)
        self.p16 = torch.zeros(self.n_shadow, dtype=torch.bfloat16, device=self.device)
        with torch.no_grad():
            for _, param, off in self.named_slots():
                view = self.p32[off : off + param.numel()].view(param.shape)
                view.copy_(param.detach().to(torch.float32))
                param.data = view
        if self.model.tie_embeddings:
            assert self.model.lm_head.weight is self.model.token_embedding.weight
        self._bind()
        self.shadow_fresh = False

    def ensure_grad_arena(self) -> torch.Tensor:
        if self.g32 is None:
            self.g32 = torch.zeros(self.n_params, dtype=torch.float32, device=self.device)
            self._bind()
        return self.g32

    def refresh_shadow(self) -> None:
        check(self._lib.llmtrain_b200_gpt_refresh_shadow(self._handle, stream_handle()),
              "gpt_refresh_shadow")
        self.shadow_fresh = True

    # ------------------------------------------------------------------ execution
    def fit_plan(self, batch: int, seq: int) -> None:
        """Grow the activation plan when a batch exceeds it (the reference accepts any batch the
        model accepts; the arena is re-laid-out, which is rare and outside the steady state)."""
        if batch > self._plan[0] or seq > self._plan[1]:
            self._replan(max(batch, self._plan[0]), max(seq, self._plan[1]))

    def _prepare(self, ids: torch.Tensor, mask: torch.Tensor | None):
        b, t = ids.shape
        self.fit_plan(b, t)
        if not self.shadow_fresh:
            self.refresh_shadow()
        ids = ids.to(self.device, non_blocking=True).contiguous()
        key_mask = None
        n_valid = b * t
        if mask is not None:
            n_valid = int((mask != 0).sum().item())  # host tensors: no device sync
            if n_valid != b * t:
                key_mask = (mask != 0).to(torch.uint8).to(self.device, non_blocking=True)
                key_mask = key_mask.contiguous()
        return ids, key_mask, n_valid

    def run_forward(
        self,
        ids: torch.Tensor,
        labels: torch.Tensor | None,
        key_mask: torch.Tensor | None,
        *,
        mode: int,
        loss_acc: torch.Tensor | None = None,
        loss_scale: float = 1.0,
        grad_scale: float = 1.0,
    ) -> None:
        b, t = ids.shape
        self._live = (ids, labels, key_mask)  # pointers are re-read by backward
        self._arm_dropout(mode)
        check(
            self._lib.llmtrain_b200_gpt_forward(
                self._handle, ptr(ids), ptr(labels), ptr(key_mask), ctypes.c_int(b),
                ctypes.c_int(t), ctypes.c_int(mode), ptr(loss_acc), ctypes.c_float(loss_scale),
                ctypes.c_float(grad_scale), stream_handle(),
            ),
            "gpt_forward",
        )

    def _arm_dropout(self, mode: int) -> None:
        """Dropout applies to training forwards (mode 2) of a module in ``train()`` mode.

        Reference: four ``nn.Dropout`` modules, ``src/llmtrain/models/gpt.py:28,30,97,129``.  Masks
        are counter-based (``csrc/common/philox.cuh``): the seed is torch's global seed plus the
        rank, the offset is a micro-batch counter, so a run is reproducible and the backward pass
        regenerates what the forward pass used.  The train engine derives the counter from the
        optimiser step (``B200TrainEngine.train_step``), which is in every checkpoint, and the
        seed from torch's generator state, which ``Trainer.restore`` puts back: a resumed run
        continues the mask stream of the interrupted one instead of replaying it from the start.
        """
        p = float(self.model.dropout_p) if (mode == 2 and self.model.training) else 0.0
        if p <= 0.0 and not self._dropout_armed:
            return
        if self._dropout_seed is None:
            rank = 0
            if torch.distributed.is_available() and torch.distributed.is_initialized():
                rank = torch.distributed.get_rank()
            self._dropout_seed = (int(torch.initial_seed()) + 0x9E3779B97F4A7C15 * rank) % (1 << 64)
        if p > 0.0:
            self._dropout_calls += 1
        check(
            self._lib.llmtrain_b200_gpt_set_dropout(
                self._handle, ctypes.c_float(p), ctypes.c_uint64(self._dropout_seed),
                ctypes.c_uint64(self._dropout_calls),
            ),
            "gpt_set_dropout",
        )
        self._dropout_armed = p > 0.0

    def run_backward(self) -> None:
        check(self._lib.llmtrain_b200_gpt_backward(self._handle, stream_handle()), "gpt_backward")

    def logits(self, input_ids: torch.Tensor, attention_mask: torch.Tensor | None) -> torch.Tensor:
        """fp32 ``(B, T, V)`` logits computed by the native kernels (inference path)."""
        ids, key_mask, _ = self._prepare(input_ids, attention_mask)
        b, t = ids.shape
        self.run_forward(ids, None, key_mask, mode=0)
        assert self._act is not None
        n = b * t * self.padded_vocab
        flat = self._act[self._logits_off : self._logits_off + 2 * n].view(torch.bfloat16)
        return flat.view(b, t, self.padded_vocab)[:, :, : self.model.vocab_size].float()

    def loss(
        self, input_ids: torch.Tensor, labels: torch.Tensor, attention_mask: torch.Tensor | None
    ) -> torch.Tensor:
        """Mean next-token loss with an autograd edge to every parameter (public-API path)."""
        ids, key_mask, n_valid = self._prepare(input_ids, attention_mask)
        if n_valid == 0:
            raise ValueError("attention_mask has no valid target tokens.")
        labels = labels.to(self.device, non_blocking=True).contiguous()
        params = [p for _, p, _ in self.named_slots()]
        needs_grad = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        return _WholeModelLoss.apply(self, ids, labels, key_mask, n_valid, needs_grad, *params)


class _WholeModelLoss(torch.autograd.Function):
    """loss = f(all parameters): forward+backward are one fused native pass.

    When gradients are wanted they are produced eagerly into a private fp32 arena during
    ``forward`` and handed to autograd (scaled by the incoming grad) in ``backward``; this keeps
    ``loss.backward()`` / ``(loss / k).backward()`` semantics, including accumulation into
    ``param.grad``, without a graph over individual ops.
    """

    @staticmethod
    def forward(ctx, engine, ids, labels, key_mask, n_valid, needs_grad, *params):  # noqa: ANN001
        acc = torch.zeros(1, dtype=torch.float32, device=engine.device)
        scale = 1.0 / n_valid
        ctx.engine = engine
        ctx.needs_grad = needs_grad
        if not needs_grad:
            engine.run_forward(ids, labels, key_mask, mode=1, loss_acc=acc, loss_scale=scale)
            return acc.reshape(())
        main = engine.g32
        private = torch.zeros(engine.n_params, dtype=torch.float32, device=engine.device)
        engine.g32 = private
        engine._bind()
        try:
            engine.run_forward(ids, labels, key_mask, mode=2, loss_acc=acc, loss_scale=scale,
                               grad_scale=scale)
            engine.run_backward()
        finally:
            engine.g32 = main
            engine._bind()
        ctx.grads = private
        return acc.reshape(())

    @staticmethod
    def backward(ctx, grad_out):  # noqa: ANN001
        engine = ctx.engine
        grads: list[torch.Tensor | None] = []
        for _, param, off in engine.named_slots():
            if not param.requires_grad:
                grads.append(None)
                continue
            g = ctx.grads[off : off + param.numel()].view(param.shape)
            grads.append(g * grad_out)
        return (None, None, None, None, None, None, *grads)


# ================================================================================================
class B200TrainEngine(TrainEngine):
    """Optimiser step on the native runtime (see module docstring)."""

    def __init__(
        self,
        model: nn.Module,
        cfg: RunConfig,
        *,
        device: torch.device,
        ddp_state: DDPState | None,
    ) -> None:
        self._cfg = cfg
        self._device = device
        self.model = model
        # trainer.extra.deterministic_kernels: fixed summation orders in the native kernels (see
        # llmtrain_b200_set_deterministic) so that a resumed run reproduces the uninterrupted one
        # bit for bit, as the reference's CPU path does under use_deterministic_algorithms
        # (`tests/test_checkpoint.py:301-320`).  Off by default: it costs throughput (weight
        # gradients lose their split-K parallelism).  `run.deterministic` keeps the reference's
        # meaning (fixed data order).
        self.deterministic = bool(cfg.trainer.extra.get("deterministic_kernels", False))
        native.lib().llmtrain_b200_set_deterministic(ctypes.c_int(1 if self.deterministic else 0))
        self._world = ddp_state.world_size if ddp_state is not None else 1
        self._rank = ddp_state.rank if ddp_state is not None else 0
        t = int(cfg.data.extra.get("seq_len", cfg.model.block_size))
        self._gpt: GPTEngine = model.native_engine(
            max_batch=cfg.trainer.micro_batch_size, max_seq=min(t, cfg.model.block_size)
        )
        g = self._gpt
        g.ensure_grad_arena()
        self._m = torch.zeros(g.n_params, dtype=torch.float32, device=device)
        self._v = torch.zeros(g.n_params, dtype=torch.float32, device=device)
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=device)
        self._norm = torch.zeros(1, dtype=torch.float32, device=device)
        self._loss_acc = torch.zeros(1, dtype=torch.float32, device=device)
        self._eval_acc = torch.zeros(1, dtype=torch.float32, device=device)
        self._step = 0
        self._comm: Any | None = None
        self._prefetch = bool(cfg.trainer.extra.get("prefetch", True))
        self._prefetched: tuple[Any, ...] | None = None
        self._last_seq_len = 0
        # Optional CUDA-event phase timeline (LLMTRAIN_PHASE_TIMING=1): forward / backward(+bucket
        # reductions) / reduce+optimiser, averaged per step in ``phase_ms``.
        import os

        self._phase_timing = os.environ.get("LLMTRAIN_PHASE_TIMING", "0") == "1"
        self._phase_events: list[list[torch.cuda.Event]] = []
        self.phase_ms: dict[str, float] = {}

        if self._world > 1:
            self._comm = self._build_comm(cfg)
            self._sync_initial_parameters()

        # torch's own AdamW object is the checkpoint-format state container; its per-parameter
        # state tensors are views of the flat moment arenas the fused kernel updates.
        self.optimizer = torch.optim.AdamW(
            model.parameters(), lr=cfg.trainer.lr, weight_decay=cfg.trainer.weight_decay
        )
        self._install_state_views()
        g.refresh_shadow()

    # ------------------------------------------------------------------ set-up helpers
    def _install_state_views(self) -> None:
        for _, param, off in self._gpt.named_slots():
            n = param.numel()
            self.optimizer.state[param] = {
                "step": torch.tensor(float(self._step), dtype=torch.float32),
                "exp_avg": self._m[off : off + n].view(param.shape),
                "exp_avg_sq": self._v[off : off + n].view(param.shape),
            }

    def _sync_initial_parameters(self) -> None:
        """Replicas must start identical: broadcast rank 0's master weights (reference: the DDP
        constructor's ``_sync_module_states``)."""
        import torch.distributed as dist

        assert self._gpt.p32 is not None
        nvl_comm = getattr(self._comm, "_comm", None)
        if nvl_comm is not None and hasattr(nvl_comm, "broadcast"):
            nvl_comm.broadcast(self._gpt.p32, root=0)  # copy engines over mapped peer memory
        else:
            dist.broadcast(self._gpt.p32, src=0)

    def _build_comm(self, cfg: RunConfig) -> Any:
        from llmtrain.parallel import nvl

        want = cfg.ddp.backend
        algo = str(cfg.trainer.extra.get("allreduce", "auto"))
        if want == "nvl" or (want == "nccl" and algo == "nvl"):
            comm = nvl.try_create(self._gpt, self._m, self._v, rank=self._rank, world=self._world,
                                  device=self._device, timeout_s=cfg.ddp.timeout_sec,
                                  overlap=bool(cfg.trainer.extra.get("overlap_comm", True)),
                                  allow_nvls=not self.deterministic)
            if comm is not None:
                return comm
            logger.warning("NVLink peer access unavailable; falling back to NCCL all-reduce")
        return nvl.NcclFallback(self._gpt, overlap=bool(cfg.trainer.extra.get("overlap_comm", True)))

    # ------------------------------------------------------------------ TrainEngine API
    @property
    def raw_model(self) -> nn.Module:
        return self.model

    def _stage(self, batch: Batch) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor | None, int]:
        ids_h, labels_h, mask_h = batch["input_ids"], batch["labels"], batch.get("attention_mask")
        b, t = ids_h.shape
        n_valid = b * t
        key_mask = None
        if mask_h is not None:
            n_valid = int(mask_h.sum().item())  # host tensor: no device sync
            if n_valid == 0:
                raise ValueError("attention_mask has no valid target tokens.")
            if n_valid != b * t:
                key_mask = (mask_h != 0).to(torch.uint8).to(self._device, non_blocking=True)
        # torch raises for a target outside [0, V) other than ignore_index (-100); the native loss
        # kernel cannot raise, so the (host-resident) labels are checked here
        if labels_h.numel() and not labels_h.is_cuda:  # device-generated batches are in range
            lo, hi = int(labels_h.min()), int(labels_h.max())
            if hi >= self.model.vocab_size or (lo < 0 and lo != -100):
                raise ValueError(
                    f"labels must lie in [0, {self.model.vocab_size}) or equal -100 (ignored); "
                    f"got min={lo} max={hi}"
                )
        ids = ids_h.to(self._device, non_blocking=True)
        labels = labels_h.to(self._device, non_blocking=True)
        return ids, labels, key_mask, n_valid

    def train_step(self, next_batch: Callable[[], Batch], accum_steps: int) -> tuple[Any, int]:
        g = self._gpt
        self._loss_acc.zero_()
        tokens = 0
        for micro in range(accum_steps):
            # The first micro-batch of a step was usually staged (pinned host -> device copy
            # enqueued) at the end of the previous step, so the copy and the loader's host work
            # overlap the optimiser kernels instead of following a loss read-back.
            staged = self._prefetched if self._prefetched is not None else self._stage(next_batch())
            self._prefetched = None
            ids, labels, key_mask, n_valid = staged
            g.fit_plan(int(ids.shape[0]), int(ids.shape[1]))
            # dropout offset = number of training micro-batches before this one: a pure function
            # of (optimiser step, position in the accumulation window), both durable
            g._dropout_calls = self._step * accum_steps + micro
            tokens += ids.numel()
            self._last_seq_len = int(ids.shape[1])
            scale = 1.0 / (n_valid * accum_steps)
            marks = self._mark()
            g.run_forward(ids, labels, key_mask, mode=2, loss_acc=self._loss_acc,
                          loss_scale=scale, grad_scale=scale)
            self._mark(marks)
            if micro == accum_steps - 1 and hasattr(self._comm, "backward_last"):
                self._comm.backward_last()  # gradients are final: reduce buckets as they finish
            else:
                g.run_backward()
            self._mark(marks)
        self._step += 1
        # the fused kernel is this optimiser's step(): tell torch's LR scheduler so, or it warns
        # about scheduler.step() preceding optimizer.step() on every run
        self.optimizer._opt_called = True  # type: ignore[attr-defined]
        lr = float(self.optimizer.param_groups[0]["lr"])
        group = self.optimizer.param_groups[0]
        hyper = dict(
            lr=lr, beta1=float(group["betas"][0]), beta2=float(group["betas"][1]),
            eps=float(group["eps"]), weight_decay=float(group["weight_decay"]), step=self._step,
            max_norm=float(self._cfg.trainer.max_grad_norm),
        )
        if self._comm is None:
            assert g.g32 is not None and g.p32 is not None
            K.grad_sumsq(g.g32, self._sumsq)
            K.adamw(g.p32, self._m, self._v, g.g32, g.p16, sumsq=self._sumsq, norm_out=self._norm,
                    **hyper)
        else:
            self._comm.reduce_and_step(self._m, self._v, self._sumsq, self._norm, hyper)
        self._mark(marks, final=True)
        loss = self._publish_scalars()
        if self._prefetch:
            self._prefetched = self._stage(next_batch())
        return loss, tokens

    _SCALAR_RING = 512

    def _publish_scalars(self) -> StepScalars:
        """Device -> pinned-host copy of (loss, grad norm) of the step just enqueued."""
        ring = getattr(self, "_scalar_ring", None)
        if ring is None:
            ring = torch.zeros(self._SCALAR_RING, 2, dtype=torch.float32).pin_memory()
            self._scalar_ring = ring
            self._scalar_events: list[torch.cuda.Event | None] = [None] * self._SCALAR_RING
            self._scalar_pending: list[StepScalars | None] = [None] * self._SCALAR_RING
        slot = self._step % self._SCALAR_RING
        stale = self._scalar_pending[slot]
        if stale is not None:
            stale._read()  # 512 steps old: complete long ago; keeps its value before the overwrite
        row = ring[slot]
        row[0:1].copy_(self._loss_acc.reshape(1), non_blocking=True)
        row[1:2].copy_(self._norm.reshape(1), non_blocking=True)
        event = torch.cuda.Event(enable_timing=_TRACE_FIT)
        event.record()
        out = StepScalars(row, event)
        if _TRACE_FIT:
            out.timing_event = event  # type: ignore[attr-defined]
        self._scalar_pending[slot] = out
        self._last_scalars = out
        return out

    def _mark(self, marks: list[torch.cuda.Event] | None = None, *, final: bool = False):
        if not self._phase_timing:
            return None
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        if marks is None:
            return [ev]
        marks.append(ev)
        if final:
            self._phase_events.append(marks)
        return marks

    def collect_phase_timing(self, skip: int = 0) -> dict[str, float]:
        """Average milliseconds per step of (forward, backward, reduce+optimiser), last micro-batch."""
        torch.cuda.synchronize()
        rows = self._phase_events[skip:]
        if rows:
            names = ("forward_ms", "backward_ms", "reduce_optim_ms")
            self.phase_ms = {
                n: sum(r[i].elapsed_time(r[i + 1]) for r in rows) / len(rows)
                for i, n in enumerate(names)
            }
        self._phase_events = []
        return self.phase_ms

    def eval_loss(self, batch: Batch) -> float:
        ids, labels, key_mask, n_valid = self._stage(batch)
        self._gpt.fit_plan(int(ids.shape[0]), int(ids.shape[1]))  # val batches may be larger
        acc = self._eval_acc
        acc.zero_()
        self._gpt.run_forward(ids, labels, key_mask, mode=1, loss_acc=acc, loss_scale=1.0 / n_valid)
        return float(acc.item())

    def before_checkpoint(self) -> None:
        for state in self.optimizer.state.values():
            state["step"] = torch.tensor(float(self._step), dtype=torch.float32)

    def after_restore(self) -> None:
        """``load_state_dict`` replaced the optimiser's state tensors with copies: pull them back
        into the flat arenas, re-alias, and rebuild the bf16 shadow from the restored weights."""
        steps = []
        with torch.no_grad():
            for _, param, off in self._gpt.named_slots():
                state = self.optimizer.state.get(param)
                if not state:
                    continue
                n = param.numel()
                self._m[off : off + n].view(param.shape).copy_(state["exp_avg"])
                self._v[off : off + n].view(param.shape).copy_(state["exp_avg_sq"])
                steps.append(int(float(state["step"])))
        self._step = max(steps) if steps else 0
        self._gpt._dropout_seed = None  # re-derive from the restored torch generator state
        self._install_state_views()
        self._gpt.refresh_shadow()

    def peak_memory_bytes(self) -> float:
        return float(torch.cuda.max_memory_allocated(self._device))

    def model_flops_per_token(self, seq_len: int) -> float:
        """Training FLOPs per token: 6 per matmul parameter (the tied table counts once, as the LM
        head; the position table does no matmul) plus 12 * L * d * T for attention, the usual MFU
        accounting (non-causal count)."""
        m = self.model
        n_matmul = sum(p.numel() for p in m.parameters()) - m.position_embedding.weight.numel()
        if not m.tie_embeddings:
            n_matmul -= m.token_embedding.weight.numel()
        return 6.0 * n_matmul + 12.0 * m.n_layers * m.d_model * seq_len

    @staticmethod
    def _peak_tflops() -> float:
        """Measured bf16 GEMM burst peak when the pod's MEASURED_PEAKS.json is present, else the
        nominal dense B200 figure."""
        import json
        from pathlib import Path

        for parent in Path(__file__).resolve().parents:
            candidate = parent / "MEASURED_PEAKS.json"
            if candidate.exists():
                try:
                    return float(json.loads(candidate.read_text())["bf16_tflops"])
                except (OSError, ValueError, KeyError):
                    break
        return 2250.0

    def extra_metrics(self, tokens_per_sec: float | None = None,
                      step_result: Any = None) -> dict[str, float]:
        # the norm of the step being logged (already on the host), not of the newest enqueued one
        norm = getattr(step_result, "grad_norm", None)
        out = {"train/grad_norm": float(self._norm.item()) if norm is None else float(norm)}
        if tokens_per_sec and self._last_seq_len:
            tflops = tokens_per_sec * self.model_flops_per_token(self._last_seq_len) / 1e12
            out["train/model_tflops_per_gpu"] = tflops
            out["train/mfu"] = tflops / self._peak_tflops()
        return out
