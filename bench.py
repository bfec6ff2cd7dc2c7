#!/usr/bin/env python
"""Headline benchmark: GPT-2 124M data-parallel training throughput (tokens/s) on N B200s.

Metric and config come from BASELINE.json: "GPT-2 124M DDP tok/s at 1/2/4/8 B200 (max over
ranks)", bf16 compute, sequence length 1024, synthetic tokens, random-init weights (there is no
network for datasets or checkpoints).  Weak scaling: the per-GPU micro-batch is fixed.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3
    python bench.py --impl reference ...        # the UNMODIFIED reference from baseline/_ref

Two numbers are produced per run, both over EXACTLY ``--steps`` optimiser steps after
``--warmup`` untimed ones, bracketed by barrier + ``torch.cuda.synchronize()`` on both sides, timed
with CUDA events on the launching stream, max over ranks:

* ``value``  — device-resident inputs, loss left on the device (what the kernels + comm can do);
* ``e2e``    — the public API (``Trainer.fit``): every step copies its inputs from pinned host
  memory and reads the step's loss back to the host.  Measured FIRST (right after its own warm-up,
  like the reference arm's single measurement): the second arm of a process runs ~1% slower
  because the board is still settling under the power cap.

For N > 1 a third, separate segment measures the exposed communication (the second half of the
BASELINE metric): the same steps with the gradient reduction switched off, in the same process;
``comm_exposed.exposed_comm_ms_per_step`` = step time with the bucket engine - step time without.

Rank 0 prints ONE JSON line.  Clocks / throttle reasons are sampled (NVML) during the timed
regions.  Inputs are far larger than L2 (weights + optimiser state alone are ~2 GB/step of
traffic), which is recorded in ``config``.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path
from typing import Any

ROOT = Path(__file__).resolve().parent
METRIC = "GPT-2 124M DDP tok/s (max over ranks)"

GPT2_SMALL = dict(block_size=1024, d_model=768, n_layers=12, n_heads=12, d_ff=3072, dropout=0.0,
                  tie_embeddings=True, vocab_size=50257)

GPT2_MEDIUM = dict(GPT2_SMALL, d_model=1024, n_layers=24, n_heads=16, d_ff=4096)
GPT2_XL = dict(GPT2_SMALL, block_size=2048, d_model=1600, n_layers=48, n_heads=25, d_ff=6400)
MODELS = {"small": GPT2_SMALL, "medium": GPT2_MEDIUM, "xl": GPT2_XL}
MODEL_NAMES = {
    "small": "gpt2-small-124M (d768 L12 H12 F3072 V50257 tied)",
    "medium": "gpt2-medium-350M (d1024 L24 H16 F4096 V50257 tied)",
    "xl": "gpt2-xl-1.5B (d1600 L48 H25 F6400 V50257 tied)",
}


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU every 100 ms on a background thread."""

    _REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap",
        0x8: "hw_slowdown", 0x10: "sync_boost", 0x20: "sw_thermal_slowdown",
        0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index: int) -> None:
        self._index = index
        self._stop = threading.Event()
        self._thread: threading.Thread | None = None
        self.samples: list[int] = []
        self.power_w: list[float] = []
        self.reasons: set[str] = set()
        self.max_mhz: int | None = None
        self.error: str | None = None

    def _run(self) -> None:
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            physical = self._index
            if visible:
                ids = [v.strip() for v in visible.split(",") if v.strip()]
                if self._index < len(ids) and ids[self._index].isdigit():
                    physical = int(ids[self._index])
            handle = pynvml.nvmlDeviceGetHandleByIndex(physical)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(handle, pynvml.NVML_CLOCK_SM))
            while not self._stop.is_set():
                self.samples.append(int(pynvml.nvmlDeviceGetClockInfo(handle, pynvml.NVML_CLOCK_SM)))
                self.power_w.append(pynvml.nvmlDeviceGetPowerUsage(handle) / 1000.0)
                mask = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(handle))
                for bit, name in self._REASONS.items():
                    if mask & bit and name != "gpu_idle":
                        self.reasons.add(name)
                self._stop.wait(0.1)
        except Exception as exc:  # pragma: no cover - depends on the box
            self.error = f"{type(exc).__name__}: {exc}"

    def __enter__(self) -> ClockSampler:
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc: Any) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=5)

    def summary(self) -> dict[str, Any]:
        out: dict[str, Any] = {
            "sm_mhz": statistics.median(self.samples) if self.samples else None,
            "sm_max_mhz": self.max_mhz,
            "reasons": sorted(self.reasons),
            "samples": len(self.samples),
            "power_w": round(statistics.median(self.power_w), 1) if self.power_w else None,
        }
        if self.error:
            out["error"] = self.error
        return out


# ------------------------------------------------------------------------------------------------
# shared helpers
# ------------------------------------------------------------------------------------------------
def _dist_env() -> tuple[int, int, int]:
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def _max_over_ranks(ms: float, world: int) -> float:
    if world == 1:
        return ms
    import torch
    import torch.distributed as dist

    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _gather_rows(row: list[float], world: int) -> list[list[float]]:
    """Every rank's row of floats, in rank order (NCCL all-gather of a small float64 tensor)."""
    if world == 1:
        return [row]
    import torch
    import torch.distributed as dist

    mine = torch.tensor(row, dtype=torch.float64, device="cuda")
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine)
    return [[float(x) for x in t.tolist()] for t in out]


def _barrier(world: int) -> None:
    import torch

    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    torch.cuda.synchronize()


def _timed(fn: Any, world: int) -> float:
    """Run ``fn`` between barrier+sync fences; return CUDA-event milliseconds (this rank)."""
    import torch

    _barrier(world)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    fn()
    stop.record()
    _barrier(world)
    return float(start.elapsed_time(stop))


def _payload(cfg_model: dict[str, Any], micro_batch: int, accum: int, world: int, seq: int,
             steps: int) -> dict[str, Any]:
    return {
        "run": {"name": "bench-gpt2-124m", "device": "cuda", "seed": 1234},
        "model": {"name": "gpt", **cfg_model},
        "data": {"name": "synthetic_text", "num_workers": 0,
                 "extra": {"num_examples": max(256, 2 * micro_batch * accum), "seq_len": seq}},
        "trainer": {"max_steps": max(steps, 1), "micro_batch_size": micro_batch,
                    "grad_accum_steps": accum, "lr": 3e-4, "weight_decay": 0.1, "warmup_steps": 0,
                    "max_grad_norm": 1.0, "log_every_steps": 1, "eval_every_steps": 10**9,
                    "save_every_steps": 10**9,
                    "extra": {"overlap_comm": os.environ.get("LLMTRAIN_OVERLAP", "1") != "0",
                              # cost measurement of the fixed-order kernels, not the headline
                              "deterministic_kernels":
                                  os.environ.get("LLMTRAIN_DETERMINISTIC", "0") == "1"}},
        # LLMTRAIN_BENCH_BACKEND=nccl measures the fallback transport (pods without shared IPC)
        "ddp": {"enabled": world > 1, "backend": os.environ.get("LLMTRAIN_BENCH_BACKEND", "nvl")},
        "mlflow": {"enabled": False},
        "logging": {"level": "WARNING", "log_to_file": False},
        "output": {"root_dir": "runs"},
    }


def metric_name(model: str) -> str:
    """BASELINE.json's headline string for the default model; the same sentence with the model's
    name for configs #3 / #4."""
    return METRIC if model == "small" else METRIC.replace(
        "GPT-2 124M", MODEL_NAMES[model].split(" ")[0])


def bench_config(mb: int, accum: int, seq: int, world: int, model: str = "small") -> dict[str, Any]:
    """The ``config`` object of the JSON line: identical for both arms by construction (the driver
    compares the two dictionaries); anything implementation-specific goes under ``notes``."""
    return {
        "model": MODEL_NAMES[model],
        "global_batch": mb * accum * world,
        "micro_batch_per_gpu": mb,
        "grad_accum": accum,
        "seq_len": seq,
        "parallelism": f"dp{world}",
        "optimizer": "AdamW on fp32 parameters + clip 1.0 (inside the timed region)",
        "l2": "inputs larger than L2 (>=2 GB of weights/optimizer state + GBs of activations "
              "touched per step vs 126 MB L2); no explicit flush",
    }


def _ddp_check(engine: Any, batch: dict[str, Any], world: int) -> dict[str, Any]:
    """Multi-GPU correctness evidence inside the bench run itself (the driver's pytest box has one
    GPU).  (1) replicas bit-identical: two integer checksums of the fp32 master arena, gathered.
    (2) the gradient the live reduction path (P2P buckets at 2 GPUs, in-switch buckets from 4)
    produces equals the sum of the ranks' local gradients of the same micro-batch, summed by NCCL
    in fp32 — to the tolerance of the wire format (bf16 for P2P, fp32 in-switch)."""
    import torch
    import torch.distributed as dist

    g, comm = engine._gpt, engine._comm
    bits = g.p32.view(torch.int32).to(torch.int64)
    weights = (torch.arange(bits.numel(), device="cuda", dtype=torch.int64) % 65521) + 1
    sums = _gather_rows([float(bits.sum().item() % (1 << 52)),
                         float((bits * weights).sum().item() % (1 << 52))], world)
    identical = all(r == sums[0] for r in sums)
    del bits, weights

    ids, labels = batch["input_ids"], batch["labels"]
    scale = 1.0 / ids.numel()

    def forward() -> None:
        engine._loss_acc.zero_()
        g.run_forward(ids, labels, None, mode=2, loss_acc=engine._loss_acc, loss_scale=scale,
                      grad_scale=scale)

    g.g32.zero_()
    forward()
    g.run_backward()  # local gradient only
    torch.cuda.synchronize()
    want = g.g32.clone()
    dist.all_reduce(want, op=dist.ReduceOp.SUM)
    g.g32.zero_()
    forward()
    reduced_by = "none"
    if hasattr(comm, "backward_last") and getattr(comm, "overlap", False):
        comm.backward_last()  # the bucket engine: reduced while backward runs
        reduced_by = "bucket engine (overlapped with backward)"
    else:
        g.run_backward()
    torch.cuda.synchronize()
    dist.barrier()  # peers' pushes into this rank's buffers have landed
    torch.cuda.synchronize()
    got = comm.reduced_gradient() if reduced_by != "none" else g.g32
    denom = float(want.abs().max().item()) + 1e-30
    err = float((got - want).abs().max().item()) / denom
    rms = float((got - want).pow(2).mean().sqrt().item()) / (
        float(want.pow(2).mean().sqrt().item()) + 1e-30)
    worst = max(r[0] for r in _gather_rows([err], world))
    g.g32.zero_()
    if hasattr(comm, "_reduced"):
        comm._reduced = False
    if hasattr(comm, "_works"):
        comm._works.clear()
    torch.cuda.synchronize()
    dist.barrier()
    tol = 2e-2 if reduced_by != "none" else 0.0
    out = {
        "replicas_bit_identical": identical,
        "path": getattr(comm, "name", "?"),
        "reduced_by": reduced_by,
        "grad_max_abs_err_over_max_abs": worst,
        "grad_rms_err_over_rms_rank0": rms,
        "tolerance": tol,
    }
    out["ok"] = bool(identical and (reduced_by == "none" or worst <= tol))
    return out


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def run_ours(args: argparse.Namespace) -> dict[str, Any]:
    sys.path.insert(0, str(ROOT / "src"))
    import torch

    from llmtrain.config.schemas import RunConfig
    from llmtrain.distributed import setup_ddp, teardown_ddp
    from llmtrain.engine.b200 import B200TrainEngine
    from llmtrain.ops import native
    from llmtrain.training import Trainer

    rank, local_rank, world = _dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun for N>1")
    seq, mb, accum = args.seq_len, args.micro_batch, args.grad_accum
    spec = MODELS[args.model]
    cfg = RunConfig.model_validate(_payload(spec, mb, accum, world, seq, args.steps + args.warmup))
    ddp_state = setup_ddp(cfg) if world > 1 else None
    torch.cuda.set_device(local_rank % torch.cuda.device_count())
    try:
        trainer = Trainer(cfg, ddp_state=ddp_state)
        engine = trainer._engine
        if not isinstance(engine, B200TrainEngine):
            raise SystemExit("bench: the native B200 engine was not selected")
        tokens_per_step = mb * accum * seq * world
        lib = native.lib()
        lib.llmtrain_b200_launch_count.restype = __import__("ctypes").c_longlong

        def run_e2e() -> tuple[float, Any]:
            trainer.fit(max_steps_override=args.warmup)
            with ClockSampler(torch.cuda.current_device()) as sampled:
                ms_ = _timed(lambda: trainer.fit(max_steps_override=args.steps), world)
            return _max_over_ranks(ms_, world), sampled

        # Which arm runs first matters by ~1%: under the power cap the board settles to a lower
        # clock over the first seconds of load (measured, same box, K = 20: device-timed 32.12 /
        # 32.08 then e2e 32.34 / 32.47 ms; e2e 32.06 / 32.16 then device-timed 32.16 / 32.45 ms).
        # The end-to-end arm goes first, as the reference arm's only measurement does;
        # LLMTRAIN_BENCH_E2E_FIRST=0 swaps the order.
        e2e_first = os.environ.get("LLMTRAIN_BENCH_E2E_FIRST", "1") == "1"
        if e2e_first:
            ms_e2e, clocks_e2e = run_e2e()

        # ---- device-resident arm ---------------------------------------------------------------
        # (tokens come from the library's device-resident generator, SURVEY W13: each rank draws
        # its own rows of one global counter-based stream)
        from llmtrain.ops import kernels as K

        pool = []
        for i in range(4):
            ids_, labels_ = K.synthetic_tokens(mb, seq, spec["vocab_size"], seed=1234,
                                               first_row=(i * world + rank) * mb)
            pool.append({"input_ids": ids_, "labels": labels_})
        cursor = [0]

        def next_device_batch() -> dict[str, Any]:
            cursor[0] += 1
            return pool[cursor[0] % len(pool)]

        step_marks: list[Any] = []

        def steps_device(n: int, *, mark: bool = False) -> None:
            for _ in range(n):
                engine.train_step(next_device_batch, accum)
                if mark:
                    ev = torch.cuda.Event(enable_timing=True)
                    ev.record()
                    step_marks.append(ev)

        steps_device(args.warmup)
        engine.collect_phase_timing()  # drop warm-up samples
        launches_before = int(lib.llmtrain_b200_launch_count())
        with ClockSampler(torch.cuda.current_device()) as clocks:
            ms = _timed(lambda: steps_device(args.steps, mark=True), world)
        launches = int(lib.llmtrain_b200_launch_count()) - launches_before
        ms_rank = ms
        ms = _max_over_ranks(ms, world)
        # per-rank attribution of the N-GPU step time: this rank's total, its per-step spread
        # (event to event), its SM clock and board power during the timed region
        per_step = [step_marks[i].elapsed_time(step_marks[i + 1]) for i in range(len(step_marks) - 1)]
        clk = clocks.summary()
        rows = _gather_rows(
            [ms_rank / args.steps,
             min(per_step) if per_step else 0.0,
             statistics.median(per_step) if per_step else 0.0,
             max(per_step) if per_step else 0.0,
             float(clk["sm_mhz"] or 0), float(clk["power_w"] or 0)], world)
        per_rank = {
            "ms_per_step": [round(r[0], 3) for r in rows],
            "step_ms_min": [round(r[1], 3) for r in rows],
            "step_ms_median": [round(r[2], 3) for r in rows],
            "step_ms_max": [round(r[3], 3) for r in rows],
            "sm_mhz": [r[4] for r in rows],
            "power_w": [r[5] for r in rows],
        }
        value = tokens_per_step * args.steps / (ms / 1e3)
        phases = engine.collect_phase_timing()

        # ---- end-to-end arm: Trainer.fit, pinned-host inputs every step, loss read every step ---
        if not e2e_first:
            ms_e2e, clocks_e2e = run_e2e()
        e2e_value = tokens_per_step * args.steps / (ms_e2e / 1e3)
        h2d = accum * (2 * mb * seq * 8)  # input_ids + labels, int64 (the all-ones mask stays home)
        d2h = 4 + 4                       # loss + grad-norm scalars read at the per-step log

        comm = getattr(engine, "_comm", None)
        ddp_check = None
        if world > 1 and comm is not None:
            engine._prefetched = None
            try:
                ddp_check = _ddp_check(engine, pool[0], world)
            except Exception as exc:  # evidence only: must never take the measurement down
                ddp_check = {"ok": False, "error": f"{type(exc).__name__}: {exc}"[:300]}
        # ---- exposed communication: the same steps with the gradient reduction switched off ----
        # (every rank then applies its local gradient; the replicas diverge, nothing is timed after
        # this).  exposed = step time with the bucket engine - step time without it, both
        # device-timed, max over ranks, same process, same clocks.
        exposed = None
        if world > 1 and comm is not None:
            engine._comm = None
            try:
                steps_device(2)
                ms_local = _max_over_ranks(_timed(lambda: steps_device(args.steps), world), world)
            finally:
                engine._comm = comm
            exposed = {"compute_only_ms_per_step": ms_local / args.steps,
                       "exposed_comm_ms_per_step": (ms - ms_local) / args.steps}
        result = {
            "metric": metric_name(args.model),
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
            "impl": "ours",
            "config": bench_config(mb, accum, seq, world, args.model),
            "notes": {"comm": getattr(comm, "name", "none") if world > 1 else "none",
                      "master_weights": "fp32 arena + bf16 shadow, fused clip+AdamW kernel"},
            "clocks": clocks.summary(),
            "per_rank": per_rank,
            "e2e": {
                "value": e2e_value,
                "unit": "tokens/s",
                "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h,
                "api": "llmtrain.training.Trainer.fit (log_every_steps=1)",
                "clocks": clocks_e2e.summary(),
            },
            "gpu_launches": launches,
            "peak_memory_gb": torch.cuda.max_memory_allocated() / 2**30,
        }
        if exposed is not None:
            result["comm_exposed"] = exposed
        if ddp_check is not None:
            result["ddp_check"] = ddp_check
        if phases:
            result["phase_ms_rank0"] = phases
        return result if rank == 0 else {}
    finally:
        if ddp_state is not None:
            teardown_ddp()


# ------------------------------------------------------------------------------------------------
# reference arm: the unmodified reference package from baseline/_ref
# ------------------------------------------------------------------------------------------------
def run_reference(args: argparse.Namespace) -> dict[str, Any]:
    ref_root = ROOT / "baseline" / "_ref"
    if not (ref_root / "llmtrain" / "training" / "trainer.py").exists():
        return {"impl": "reference",
                "unavailable": "baseline/_ref is missing (pip install --target baseline/_ref failed)"}
    sys.path.insert(0, str(ref_root))
    sys.path.insert(0, str(ROOT / "baseline"))
    try:
        import ref_harness
    except Exception as exc:
        return {"impl": "reference", "unavailable": f"harness import failed: {exc}"}
    rank, local_rank, world = _dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun for N>1")
    import torch

    if not torch.cuda.is_available():
        return {"impl": "reference", "unavailable": "no CUDA device visible to this process"}
    try:
        return ref_harness.run(args, rank=rank, local_rank=local_rank, world=world,
                               model=MODELS[args.model], metric=metric_name(args.model),
                               timed=_timed,
                               max_over_ranks=_max_over_ranks, sampler=ClockSampler,
                               config=bench_config(args.micro_batch, args.grad_accum,
                                                   args.seq_len, world, args.model))
    except Exception as exc:  # the arm must never take the driver down
        if rank == 0:
            return {"impl": "reference", "unavailable": f"{type(exc).__name__}: {exc}"[:300]}
        return {}


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--micro-batch", type=int, default=32, help="sequences per GPU per micro-step")
    ap.add_argument("--grad-accum", type=int, default=1)
    ap.add_argument("--seq-len", type=int, default=1024)
    ap.add_argument("--model", choices=sorted(MODELS), default="small",
                    help="small = the headline config; medium / xl = BASELINE configs #3 / #4")
    args = ap.parse_args()
    if args.seq_len > MODELS[args.model]["block_size"]:
        raise SystemExit("--seq-len exceeds the model's block_size")
    args.warmup = max(args.warmup, 3)  # timing rule: at least 3 warm-up steps
    result = run_reference(args) if args.impl == "reference" else run_ours(args)
    if result:
        print(json.dumps(result), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
