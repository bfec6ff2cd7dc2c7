#!/usr/bin/env python
"""Parameter counts of GPT shapes — the script form of `notebooks/gpt_model_smoke.ipynb`.

    python scripts/param_count.py                       # the notebook's 118,528-parameter model
    python scripts/param_count.py --preset gpt2_small   # 124,439,808
"""

from __future__ import annotations

import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "src"))

SHAPES = {
    "notebook": dict(vocab_size=256, block_size=32, d_model=64, n_layers=2, n_heads=4, d_ff=256),
    "gpt2_small": dict(vocab_size=50257, block_size=1024, d_model=768, n_layers=12, n_heads=12, d_ff=3072),
    "gpt2_medium": dict(vocab_size=50257, block_size=1024, d_model=1024, n_layers=24, n_heads=16, d_ff=4096),
    "gpt2_xl": dict(vocab_size=50257, block_size=2048, d_model=1600, n_layers=48, n_heads=25, d_ff=6400),
}


def main() -> int:
    import torch

    from llmtrain.models.gpt import GPT

    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", choices=sorted(SHAPES), default="notebook")
    args = ap.parse_args()
    with torch.device("meta"):
        model = GPT(dropout=0.0, **SHAPES[args.preset])
    total = sum(p.numel() for p in model.parameters())
    print(f"{args.preset}: {total:,} parameters "
          f"({total * 4 / 2**20:.1f} MiB fp32, tied embeddings counted once)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
