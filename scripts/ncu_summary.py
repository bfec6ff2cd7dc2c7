#!/usr/bin/env python
"""One-screen summary of an ``.ncu-rep`` (read here, on the CPU box): per captured launch the
duration, registers, tensor-pipe / MUFU / issue utilisation, DRAM bytes, the warp-stall mix, and
the ten instructions with the most stall samples.

    python scripts/ncu_summary.py gpurun_out/attn.ncu-rep > profiles/attention_ncu_r2.md
"""

from __future__ import annotations

import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "smsp__cycles_active.avg", "smsp__inst_executed.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.per_cycle_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
]


def page(rep: str, name: str) -> str:
    return subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True,
                          text=True, check=False).stdout


def main() -> int:
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(page(rep, "raw"))))
    head, units = rows[0], rows[1]
    print(f"# ncu summary of `{rep.split('/')[-1]}` (--set full, --clock-control none)\n")
    for r in rows[2:]:
        rec = dict(zip(head, r, strict=False))
        unit = dict(zip(head, units, strict=False))
        print(f"## {rec.get('Kernel Name', '?')[:110]}\n")
        print("| metric | value |\n|---|---|")
        for k in KEYS:
            if k in rec:
                print(f"| `{k}` | {rec[k]} {unit.get(k, '')} |")
        stalls = sorted(
            ((float(v or 0), k.split("issue_stalled_")[1].split("_per_issue")[0])
             for k, v in rec.items()
             if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith(".ratio")),
            reverse=True)[:6]
        print("| warp stalls per issue (top) | "
              + ", ".join(f"{n} {v:.2f}" for v, n in stalls) + " |\n")
    src = page(rep, "source")
    for part in src.split('"Kernel Name"')[1:]:
        lines = ('"Kernel Name"' + part).splitlines()
        table = list(csv.reader(lines[1:]))
        hdr = table[0]
        data = [r for r in table[1:] if len(r) == len(hdr)]
        ix = {h: i for i, h in enumerate(hdr)}
        if "# Samples" not in ix:
            continue
        total = sum(int(r[ix["# Samples"]]) for r in data) or 1
        print(f"### hottest instructions: {lines[0].split(',')[1][:90]}  ({total} samples)\n")
        print("| samples | share | SASS | dominant stall |\n|---:|---:|---|---|")
        for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[:10]:
            n = int(r[ix["# Samples"]])
            st = max(((int(r[ix[k]]), k) for k in hdr if k.startswith("stall_") and "Not" not in k),
                     default=(0, ""))
            print(f"| {n} | {100 * n / total:.1f}% | `{r[ix['Source']].strip()[:70]}` | {st[1]} |")
        print()
    return 0


if __name__ == "__main__":
    sys.exit(main())
