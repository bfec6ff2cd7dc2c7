#!/usr/bin/env python
"""Per-kernel-class roofline table for one training step of the bench configuration.

Input: an `ncu --metrics gpu__time_duration.sum` launch list of ONE step (bench/profile_step.py).
The work per class is computed from the model shape (GPT-2 small, M = micro_batch * seq tokens);
fractions are "of measured" peaks from MEASURED_PEAKS.json (sustained bf16 GEMM, HBM copy).

    python scripts/roofline_table.py gpurun_out/launches3.csv [--micro-batch 16]
"""

from __future__ import annotations

import collections
import csv
import json
import re
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def main() -> int:
    path = sys.argv[1]
    mb = int(sys.argv[sys.argv.index("--micro-batch") + 1]) if "--micro-batch" in sys.argv else 16
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    tf_peak = float(peaks.get("bf16_tflops", 1590.0))
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    T, d, F, H, L, V, Vp = 1024, 768, 3072, 12, 12, 50257, 50304
    M = mb * T
    n_params = 124_439_808
    lines = [ln for ln in open(path) if not ln.startswith("==")]
    agg: dict[str, list[float]] = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        val = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        ms = val / 1e6 if unit in ("ns", "nsecond") else (val / 1e3 if unit.startswith("us") else val)
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("b200::", "")
        agg[name][0] += 1
        agg[name][1] += ms
    g = lambda m, n, k: 2.0 * m * n * k  # noqa: E731
    attn = 4.0 * mb * H * T * T * 64 / 2
    # out_proj's dgrad runs with the row-statistics epilogue (<0, 1, 6>) in launch lists taken after
    # that fusion; older lists have it in the plain dgrad class
    rowdot = any(k.replace("true", "1").replace("false", "0").startswith("gemm::gemm_kernel<0, 1, 6")
                 for k in agg)
    work = {  # class -> ("flop" | "byte", amount per step)
        "gemm::gemm_kernel<0, 0, 0>": ("flop", L * g(M, 3 * d, d) + g(M, Vp, d)),
        "gemm::gemm_kernel<0, 0, 1>": ("flop", L * g(M, F, d)),
        "gemm::gemm_kernel<0, 0, 2>": ("flop", L * (g(M, d, d) + g(M, d, F))),
        "gemm::gemm_kernel<0, 1, 0>": ("flop", L * (g(M, d, 3 * d) + (0 if rowdot else g(M, d, d)) + g(M, d, F)) + g(M, d, Vp)),
        "gemm::gemm_kernel<0, 1, 6>": ("flop", L * g(M, d, d)),
        "gemm::gemm_kernel<0, 1, 3>": ("flop", L * g(M, F, d)),
        "gemm::gemm_kernel<1, 1, 4>": ("flop", L * (g(3 * d, d, M) + g(d, d, M) + 2 * g(F, d, M)) + g(V, d, M)),
        "attn::fwd_kernel": ("flop", L * attn),
        "attn_bwd::bwd_kernel": ("flop", L * attn * 2.5),
        "attn_bwd_t::bwd_kernel": ("flop", L * attn * 2.5),
        "ew::ce_kernel<1>": ("byte", 2.0 * M * Vp * 2),
        "ew::ce_row_kernel<1>": ("byte", 2.0 * M * Vp * 2),
        "ln::fwd_kernel": ("byte", (2 * L + 1) * 2.0 * M * d * 2),
        "ln::bwd_kernel": ("byte", (2 * L + 1) * 4.0 * M * d * 2),
        "ew::colsum_partial_kernel": ("byte", L * (3 * d + F) * M * 2.0),
        "optim::adamw_kernel": ("byte", n_params * (16.0 + 14.0 + 4.0)),
        "attn_bwd::dsum_kernel": ("byte", L * 2.0 * M * d * 2),
        "attn_bwd_t::stats_kernel": ("byte", L * 2.0 * M * d * 2),
        "attn_bwd::dq_convert_kernel": ("byte", L * (M * d * 4.0 + M * d * 2.0)),
    }
    total = sum(v[1] for v in agg.values())
    print(f"step = {total:.2f} ms over {int(sum(v[0] for v in agg.values()))} launches, "
          f"M = {M} tokens; peaks: {tf_peak:.0f} TFLOP/s bf16 (measured burst), {hbm:.0f} GB/s HBM\n")
    print("| kernel class | launches | ms | share | achieved | of measured peak |")
    print("|---|---:|---:|---:|---:|---:|")
    for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        norm = name.replace("true", "1").replace("false", "0")
        # keys name the first template arguments only (gemm_kernel<A_MN, B_MN, EPI[, CTAS]>)
        key = next((k for k in work if name.startswith(k.split("<")[0]) and
                    (("<" not in k) or norm.startswith(k.rstrip(">")))), None)
        ach = frac = ""
        if key is not None:
            kind, amount = work[key]
            if kind == "flop":
                rate = amount / (ms / 1e3) / 1e12
                ach, frac = f"{rate:.0f} TFLOP/s", f"{rate / tf_peak:.2f}"
            else:
                rate = amount / (ms / 1e3) / 1e9
                ach, frac = f"{rate:.0f} GB/s", f"{rate / hbm:.2f}"
        print(f"| `{name}` | {int(n)} | {ms:.3f} | {100 * ms / total:.1f}% | {ach} | {frac} |")
    return 0


if __name__ == "__main__":
    sys.exit(main())
