#!/usr/bin/env python
"""Turn an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table.

    python scripts/summarize_launches.py gpurun_out/launches.csv [--md]
"""

from __future__ import annotations

import collections
import csv
import re
import sys


def main() -> int:
    path = sys.argv[1]
    as_md = "--md" in sys.argv
    lines = [ln for ln in open(path) if not ln.startswith("==")]
    agg: dict[str, list[float]] = collections.defaultdict(lambda: [0, 0.0])
    total = 0.0
    for row in csv.DictReader(lines):
        val = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        ms = val / 1e6 if unit in ("ns", "nsecond") else (val / 1e3 if unit.startswith("us") else val)
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "")
        agg[name][0] += 1
        agg[name][1] += ms
        total += ms
    print(f"total {total:.2f} ms over {int(sum(v[0] for v in agg.values()))} launches "
          "(serialised, cold caches: compare shares, not absolutes)\n")
    if as_md:
        print("| kernel | launches | ms | share |\n|---|---:|---:|---:|")
    for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if as_md:
            print(f"| `{name}` | {int(n)} | {ms:.3f} | {100 * ms / total:.1f}% |")
        else:
            print(f"{ms:8.3f} ms {100 * ms / total:5.1f}%  n={int(n):3d}  {name}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
