#!/usr/bin/env python
"""Attention forward/backward timing at the bench shape (CUDA events), ncu-friendly (few launches)."""

from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def main() -> int:
    import torch

    from llmtrain.ops import kernels as K

    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--heads", type=int, default=12)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--dropout", type=float, default=0.0,
                    help="probability dropout (takes the row-oriented backward kernel)")
    ap.add_argument("--no-dbias", action="store_true",
                    help="do not accumulate the qkv_proj bias gradient in the backward kernels")
    args = ap.parse_args()
    b, t, h = args.batch, args.seq, args.heads
    d = h * 64
    qkv = (torch.randn(b, t, 3 * d, device="cuda")).to(torch.bfloat16)
    dout = torch.randn(b, t, d, device="cuda").to(torch.bfloat16)
    out, lse = K.attention_fwd(qkv, h)
    dqkv = torch.empty_like(qkv)

    def timeit(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.iters):
            fn()
        e.record()
        torch.cuda.synchronize()
        return a.elapsed_time(e) / args.iters

    drop = (args.dropout, 1234, 0, 1) if args.dropout > 0 else None
    fwd_ms = timeit(lambda: K.attention_fwd(qkv, h, out=out, lse=lse, dropout=drop))
    # as in the training step: qkv_proj's bias gradient is accumulated by the same kernels
    dbias = None if args.no_dbias else torch.zeros(3 * h * 64, device="cuda")
    bwd_ms = timeit(lambda: K.attention_bwd(qkv, out, dout, lse, h, dqkv=dqkv, dbias=dbias,
                                            dropout=drop))
    flops_fwd = 4.0 * b * h * t * t * 64 / 2  # causal
    print(json.dumps({
        "B": b, "T": t, "H": h, "dropout": args.dropout, "fwd_ms": round(fwd_ms, 4), "bwd_ms": round(bwd_ms, 4),
        "fwd_tflops": round(flops_fwd / fwd_ms / 1e9, 1),
        "bwd_tflops": round(2.5 * flops_fwd / bwd_ms / 1e9, 1),
    }))
    return 0


if __name__ == "__main__":
    sys.exit(main())
