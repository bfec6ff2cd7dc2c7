#!/usr/bin/env python
"""Launch the fused cross-entropy kernel a few times at the bench shape (ncu target)."""

from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def main() -> int:
    import torch

    from llmtrain.ops import kernels as K

    M, V, ldv = 16384, 50257, 50304
    logits = (torch.randn(M, ldv, device="cuda") * 2).to(torch.bfloat16)
    labels = torch.randint(0, V, (M,), device="cuda")
    rows_loss = torch.empty(M, device="cuda")
    for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
        K.cross_entropy(logits, labels, vocab=V, loss_rows=rows_loss, grad_scale=1.0 / M)
    torch.cuda.synchronize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
