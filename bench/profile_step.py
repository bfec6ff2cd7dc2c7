#!/usr/bin/env python
"""Run W warm-up + K profiled optimiser steps of the bench configuration (device-resident inputs).

Meant to be wrapped by ncu with ``--profile-from-start off``: the profiled region is bracketed by
cudaProfilerStart/Stop, so only the K steady-state steps are captured.  Without ncu it prints a
per-phase CUDA-event breakdown (forward / backward / optimiser) instead.
"""

from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))
sys.path.insert(0, str(ROOT))


def main() -> int:
    import torch

    import bench as B
    from llmtrain.config.schemas import RunConfig
    from llmtrain.training import Trainer

    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--micro-batch", type=int, default=16)
    ap.add_argument("--seq-len", type=int, default=1024)
    ap.add_argument("--layers", type=int, default=12)
    args = ap.parse_args()

    model = dict(B.GPT2_SMALL, n_layers=args.layers)
    cfg = RunConfig.model_validate(B._payload(model, args.micro_batch, 1, 1, args.seq_len, 100))
    trainer = Trainer(cfg)
    engine = trainer._engine
    gpt = engine._gpt
    mb, seq = args.micro_batch, args.seq_len
    stream_ = torch.randint(0, 50257, (mb, seq + 1), device="cuda")
    batch = {"input_ids": stream_[:, :-1].contiguous(), "labels": stream_[:, 1:].contiguous()}

    for _ in range(args.warmup):
        engine.train_step(lambda: batch, 1)
    torch.cuda.synchronize()

    # phase breakdown with CUDA events (not under a profiler)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    acc = torch.zeros(1, device="cuda")
    scale = 1.0 / (mb * seq)
    phases = {"forward+ce": 0.0, "backward": 0.0, "optimizer": 0.0}
    from llmtrain.ops import kernels as K

    torch.cuda.profiler.start()
    for _ in range(args.steps):
        ev[0].record()
        gpt.run_forward(batch["input_ids"], batch["labels"], None, mode=2, loss_acc=acc,
                        loss_scale=scale, grad_scale=scale)
        ev[1].record()
        gpt.run_backward()
        ev[2].record()
        K.grad_sumsq(gpt.g32, engine._sumsq)
        K.adamw(gpt.p32, engine._m, engine._v, gpt.g32, gpt.p16, sumsq=engine._sumsq,
                norm_out=engine._norm, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8,
                weight_decay=0.1, step=10, max_norm=1.0)
        ev[3].record()
        torch.cuda.synchronize()
        phases["forward+ce"] += ev[0].elapsed_time(ev[1])
        phases["backward"] += ev[1].elapsed_time(ev[2])
        phases["optimizer"] += ev[2].elapsed_time(ev[3])
    torch.cuda.profiler.stop()
    out = {k: v / args.steps for k, v in phases.items()}
    out["total_ms"] = sum(out.values())
    out["tokens_per_s"] = mb * seq / (out["total_ms"] / 1e3)
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
