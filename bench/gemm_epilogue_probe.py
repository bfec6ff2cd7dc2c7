#!/usr/bin/env python
"""Launch each fused-epilogue GEMM variant a few times at the bench shape (ncu target)."""

from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def main() -> int:
    import torch

    from llmtrain.ops import gemm as G

    M, d, F = 16384, 768, 3072

    def bf(*shape):
        return (torch.randn(*shape, device="cuda") * 0.05).to(torch.bfloat16)

    x, wfc, bias = bf(M, d), bf(F, d), torch.randn(F, device="cuda")
    pre, act = torch.empty(M, F, device="cuda", dtype=torch.bfloat16), torch.empty(M, F, device="cuda", dtype=torch.bfloat16)
    dres, wproj = bf(M, d), bf(d, F)
    dfc = torch.empty(M, F, device="cuda", dtype=torch.bfloat16)
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3

    def t(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters * 1e3

    print("fc plain   us", round(t(lambda: G.gemm(G.NT, G.EPI_BIAS, x, wfc, act, bias=bias)), 1))
    print("fc gelu    us", round(t(lambda: G.gemm(G.NT, G.EPI_BIAS_GELU, x, wfc, act, c2=pre, bias=bias)), 1))
    print("proj dgrad us", round(t(lambda: G.gemm(G.NN, G.EPI_BIAS, dres, wproj, dfc)), 1))
    print("proj dgelu us", round(t(lambda: G.gemm(G.NN, G.EPI_DGELU, dres, wproj, dfc, r=pre)), 1))
    cs = torch.zeros(F, device="cuda")
    print("proj dgelu+colsum us",
          round(t(lambda: G.gemm_colsum(G.NN, G.EPI_DGELU, dres, wproj, dfc, cs, r=pre)), 1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
