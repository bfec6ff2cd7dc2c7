#!/usr/bin/env python
"""Launch the LayerNorm kernels a few times at the bench shape (ncu target)."""

from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def main() -> int:
    import torch

    from llmtrain.ops import kernels as K

    M, d = 16384, 768
    x = torch.randn(M, d, device="cuda").to(torch.bfloat16)
    dy = torch.randn(M, d, device="cuda").to(torch.bfloat16)
    dres = torch.randn(M, d, device="cuda").to(torch.bfloat16)
    gamma, beta = torch.rand(d, device="cuda") + 0.5, torch.randn(d, device="cuda")
    y = torch.empty_like(x)
    mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
    dg, db, dc = (torch.zeros(d, device="cuda") for _ in range(3))
    for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
        K.layernorm_fwd(x, gamma, beta, y=y, mean=mean, rstd=rstd)
        K.layernorm_bwd(dy, x, gamma, mean, rstd, dres, dg, db, dcolsum=dc)
    torch.cuda.synchronize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
