#!/usr/bin/env python
"""Time the memory-bound kernels one at a time at the bench shape (M = 16384 tokens, GPT-2 124M).

Each launch is bracketed by its own pair of CUDA events; a 256 MB write before every launch
evicts L2 (outside the bracket) so every launch streams from HBM.  Reports the median.
Prints one JSON line per kernel with achieved GB/s against the bytes the kernel MUST move.
"""

from __future__ import annotations

import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def main() -> int:
    import torch

    from llmtrain.ops import kernels as K

    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    M, d, V, ldv = 16384, 768, 50257, 50304
    dev = torch.device("cuda")
    flush_buf = torch.empty(64 << 20, dtype=torch.float32, device=dev)

    def bf(*shape, scale=1.0):
        return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        pairs = []
        for _ in range(iters):
            flush_buf.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            pairs.append((a, b))
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) * 1e3 for a, b in pairs)
        return ts[len(ts) // 2]

    def report(name, fn, nbytes):
        us = timed(fn)
        print(json.dumps({"kernel": name, "us": round(us, 1), "GBps": round(nbytes / us / 1e3, 0)}))

    x, dy = bf(M, d), bf(M, d)
    gamma, beta = torch.rand(d, device=dev) + 0.5, torch.randn(d, device=dev)
    y = torch.empty_like(x)
    mean, rstd = torch.empty(M, device=dev), torch.empty(M, device=dev)
    report("ln_fwd", lambda: K.layernorm_fwd(x, gamma, beta, y=y, mean=mean, rstd=rstd), 2 * M * d * 2)
    dres = bf(M, d)
    dg, db, dc = (torch.zeros(d, device=dev) for _ in range(3))
    report("ln_bwd_colsum",
           lambda: K.layernorm_bwd(dy, x, gamma, mean, rstd, dres, dg, db, dcolsum=dc), 4 * M * d * 2)
    report("ln_bwd", lambda: K.layernorm_bwd(dy, x, gamma, mean, rstd, dres, dg, db), 4 * M * d * 2)

    logits = bf(M, ldv, scale=2.0)
    labels = torch.randint(0, V, (M,), device=dev)
    acc = torch.zeros(1, device=dev)
    rows_loss = torch.empty(M, device=dev)
    report("ce_fwd_bwd",
           lambda: K.cross_entropy(logits, labels, vocab=V, loss_rows=rows_loss, loss_acc=acc,
                                   loss_scale=1.0 / M, grad_scale=1.0 / M), 2 * M * ldv * 2)
    big = bf(M, 3 * d)
    dbias = torch.zeros(3 * d, device=dev)
    report("bias_grad_2304", lambda: K.bias_grad(big, dbias), M * 3 * d * 2)
    return 0


if __name__ == "__main__":
    sys.exit(main())
