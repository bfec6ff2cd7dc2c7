#!/usr/bin/env python
"""Kernel timeline of a few training steps via torch.profiler (CUPTI): name, stream, start, dur.

Works for 1 rank or under torchrun; rank 0 writes `gpurun_out/timeline_rank0.csv`.
"""

from __future__ import annotations

import csv
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))
sys.path.insert(0, str(ROOT))


def main() -> int:
    import torch
    from torch.profiler import ProfilerActivity, profile

    import bench as B
    from llmtrain.config.schemas import RunConfig
    from llmtrain.distributed import setup_ddp, teardown_ddp
    from llmtrain.training import Trainer

    rank, local_rank, world = B._dist_env()
    cfg = RunConfig.model_validate(B._payload(B.GPT2_SMALL, 16, 1, world, 1024, 100))
    state = setup_ddp(cfg) if world > 1 else None
    torch.cuda.set_device(local_rank)
    trainer = Trainer(cfg, ddp_state=state)
    engine = trainer._engine
    stream_ = torch.randint(0, 50257, (16, 1025), device="cuda")
    batch = {"input_ids": stream_[:, :-1].contiguous(), "labels": stream_[:, 1:].contiguous()}
    for _ in range(3):
        engine.train_step(lambda: batch, 1)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(2):
            engine.train_step(lambda: batch, 1)
        torch.cuda.synchronize()
    if rank == 0:
        out = ROOT / "gpurun_out"
        out.mkdir(exist_ok=True)
        rows = []
        for ev in prof.events():
            if ev.device_type is not None and "CUDA" in str(ev.device_type):
                rows.append((ev.time_range.start, ev.time_range.end - ev.time_range.start,
                             getattr(ev, "device_index", -1), ev.name[:80]))
        # kineto events carry the stream in the raw trace; export it too
        prof.export_chrome_trace(str(out / "timeline_rank0.json"))
        rows.sort()
        with open(out / "timeline_rank0.csv", "w", newline="") as fh:
            w = csv.writer(fh)
            w.writerow(["start_us", "dur_us", "device", "name"])
            w.writerows(rows)
        print("events", len(rows))
    if state is not None:
        teardown_ddp()
    return 0


if __name__ == "__main__":
    sys.exit(main())
