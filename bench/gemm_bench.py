#!/usr/bin/env python
"""Per-shape throughput of the tcgen05 GEMM family vs torch.matmul (cuBLAS) on the same shapes.

Shapes are the GPT-2 small training GEMMs at M = micro_batch * seq tokens.  Timing: CUDA events,
3 warm-up + 10 timed launches per shape, operands rotated through > L2-sized pools.
"""

from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def _time(fn, iters=10, warm=3) -> float:
    import torch

    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main() -> int:
    import torch

    from llmtrain.ops import gemm as G

    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=16384)
    ap.add_argument("--d", type=int, default=768)
    ap.add_argument("--vocab", type=int, default=50304)
    ap.add_argument("--square", type=int, default=0,
                    help="also time an N^3 problem (asymptotic mainloop efficiency)")
    args = ap.parse_args()
    M, d, F, V = args.tokens, args.d, 4 * args.d, args.vocab
    peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()).get("bf16_tflops", 1649.3) \
        if (ROOT / "MEASURED_PEAKS.json").exists() else 1590.0

    def bf(*shape):
        return (torch.randn(*shape, device="cuda") * 0.05).to(torch.bfloat16)

    rows = []
    if args.square:
        n = args.square
        a, b = bf(n, n), bf(n, n)
        c = torch.empty(n, n, device="cuda", dtype=torch.bfloat16)
        flops = 2.0 * n * n * n
        t_ours = _time(lambda: G.gemm(G.NT, G.EPI_BIAS, a, b, c))
        t_ref = _time(lambda: torch.matmul(a, b.t()))
        print(json.dumps({"gemm": f"square{n}", "kind": "fwd NT", "ours_ms": round(t_ours, 4),
                          "cublas_ms": round(t_ref, 4),
                          "ours_tflops": round(flops / t_ours / 1e9, 1),
                          "cublas_tflops": round(flops / t_ref / 1e9, 1)}), flush=True)
    fwd = [("qkv", 3 * d, d), ("out", d, d), ("fc", F, d), ("proj", d, F), ("lm_head", V, d)]
    for name, N, K in fwd:
        x, w = bf(M, K), bf(N, K)
        y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        dy = bf(M, N)
        dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
        dw = torch.zeros(N, K, device="cuda")
        flops = 2.0 * M * N * K
        cases = {
            "fwd NT": (lambda: G.gemm(G.NT, G.EPI_BIAS, x, w, y), lambda: torch.matmul(x, w.t())),
            "dgrad NN": (lambda: G.gemm(G.NN, G.EPI_BIAS, dy, w, dx), lambda: torch.matmul(dy, w)),
            "wgrad TN": (lambda: G.gemm(G.TN, G.EPI_ACCUM_F32, dy, x, dw),
                         lambda: torch.matmul(dy.t(), x)),
        }
        for kind, (ours, ref) in cases.items():
            t_ours, t_ref = _time(ours), _time(ref)
            rows.append({
                "gemm": name, "kind": kind, "M": M, "N": N, "K": K,
                "ours_ms": round(t_ours, 4), "cublas_ms": round(t_ref, 4),
                "ours_tflops": round(flops / t_ours / 1e9, 1),
                "cublas_tflops": round(flops / t_ref / 1e9, 1),
                "ours_frac_of_measured_peak": round(flops / t_ours / 1e9 / peak, 3),
            })
            print(json.dumps(rows[-1]), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
