#!/usr/bin/env python
"""BASELINE configs #3 and #4 through the real CLI under torchrun.

  #3  GPT-2 medium 350M, grad-accum 4, checkpoint at step S/2, then `--resume` to S and compare the
      final loss with the uninterrupted run (needs N >= 1 GPUs).
  #4  GPT-2 XL 1.5B, seq 2048, a few steps (memory plan + large buckets).

    python bench/config_runs.py --gpus 2 --which 3
Prints one JSON line per run with tok/s taken from the trainer's own log lines.
"""

from __future__ import annotations

import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

import yaml

ROOT = Path(__file__).resolve().parents[1]


def run_cli(cfg: dict, gpus: int, run_id: str, extra: list[str], workdir: Path) -> dict:
    path = workdir / f"{run_id}.yaml"
    path.write_text(yaml.safe_dump(cfg, sort_keys=False))
    env = dict(os.environ, PYTHONPATH=str(ROOT / "src"))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}",
           "--master-addr", "127.0.0.1", "--master-port", "29577", "-m", "llmtrain", "train",
           "--config", str(path), "--run-id", run_id, "--json", *extra]
    proc = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=1500)
    if proc.returncode != 0:
        raise SystemExit(f"{run_id} failed:\n{proc.stderr[-3000:]}")
    # NCCL may print its version banner on stdout ahead of the summary document
    summary = json.loads(proc.stdout[proc.stdout.index("{"):])
    # per-rank, host-timed interval rates from the trainer's log lines; checkpoint steps are slow
    # by design (the save sits inside the interval), so report the median
    rates = sorted(float(m) for m in re.findall(r"tokens_per_sec=([0-9.]+)", proc.stderr))
    median = rates[len(rates) // 2] if rates else None
    return {"summary": summary["training"], "tokens_per_sec_per_rank_median": median,
            "log_lines": len(rates)}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--which", type=int, choices=(3, 4), default=3)
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()
    work = Path(tempfile.mkdtemp(prefix="llmtrain_cfg_"))
    preset = "gpt2_medium_b200_ddp" if args.which == 3 else "gpt2_xl_b200_ddp"
    cfg = yaml.safe_load((ROOT / "configs" / "presets" / f"{preset}.yaml").read_text())
    cfg["output"] = {"root_dir": str(work / "runs")}
    cfg["ddp"]["enabled"] = args.gpus > 1
    cfg["trainer"].update(max_steps=args.steps, save_every_steps=args.steps // 2, log_every_steps=1,
                          warmup_steps=0)
    cfg["data"]["extra"] = {"num_examples": 512, "pattern": "arith"}
    cfg["logging"] = {"log_to_file": False, "json_output": False}
    try:
        full = run_cli(cfg, args.gpus, "full", [], work)
        out = {"config": preset, "gpus": args.gpus, "steps": args.steps,
               "global_tokens_per_step": cfg["trainer"]["micro_batch_size"]
               * cfg["trainer"]["grad_accum_steps"] * cfg["model"]["block_size"] * args.gpus,
               "full": full}
        if args.which == 3:
            ckpt = work / "runs" / "full" / "checkpoints" / f"step_{args.steps // 2:06d}.pt"
            resumed = run_cli(cfg, args.gpus, "resumed", ["--resume", str(ckpt)], work)
            out["resumed"] = resumed
            out["final_loss_abs_diff"] = abs(resumed["summary"]["final_loss"]
                                            - full["summary"]["final_loss"])
        print(json.dumps(out))
    finally:
        shutil.rmtree(work, ignore_errors=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
