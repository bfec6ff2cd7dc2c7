#!/usr/bin/env python
"""Regenerates ``profiles/sass/*.txt`` from the object files of the last native build.

    python -c "import __graft_entry__ as g; g.build()" && python scripts/sass_summary.py

One file per object: ``cuobjdump -sass`` filtered to the instructions that identify the hardware
path (tcgen05, TMA, mbarrier, multimem, system-scope memory operations, atomics, MUFU), with the
per-kernel counts the documents quote.  Runs on the CPU box (no GPU needed).
"""

from __future__ import annotations

import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
KEEP = re.compile(
    r"UTC\w*MMA|UTCBAR|UTCCP|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|UTMAPF|UBLKCP|UBLKRED|SYNCS|"
    r"ACQBULK|ELECT|MEMBAR|ERRBAR|CCTL|LDGMC|REDGMC|MULTIMEM|\.SYS\b|ATOM|\bRED\b|REDG|REDUX|"
    r"MUFU|NANOSLEEP|SETMAXREG|USETMAXREG|FENCE")
HEADER = """# cuobjdump -sass build/obj/{name}.o (sm_100a), filtered to the instructions that identify the
# hardware path: tcgen05 (UTC*MMA, UTCBAR, LDTM/STTM), TMA (UTMALDG/UTMASTG/UTMAREDG, UBLKCP),
# mbarrier (SYNCS), multimem (LDGMC/REDGMC...), system-scope LD/ST, atomics/reductions, MUFU.
# One line per instruction: address, SASS.  Counts per kernel at the end of each section.
"""
LINE = re.compile(r"^\s*/\*([0-9a-f]{4,})\*/\s+(.*?);")


def mnemonic(sass: str) -> str:
    words = sass.split()
    op = words[1] if words[0].startswith("@") and len(words) > 1 else words[0]
    base = op.split(".")[0]
    return "MUFU." + op.split(".")[1] if base == "MUFU" and "." in op else base


def summarize(obj: Path) -> str:
    text = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True,
                          check=True).stdout
    out = [HEADER.format(name=obj.stem)]
    counts: collections.Counter[str] = collections.Counter()
    open_section = False

    def close() -> None:
        if open_section:
            tally = ", ".join(f"{k} x{v}" for k, v in sorted(counts.items()))
            out.append(f"  -- counts: {tally}\n" if tally else "  -- counts: (none)\n")

    for line in text.splitlines():
        if "Function :" in line:
            close()
            counts.clear()
            open_section = True
            out.append(f"== {line.split('Function :')[1].strip()}")
            continue
        m = LINE.match(line)
        if m and open_section and KEEP.search(m.group(2)):
            sass = " ".join(m.group(2).split())
            out.append(f"  {m.group(1)}  {sass} ;")
            counts[mnemonic(sass)] += 1
    close()
    return "\n".join(out) + "\n"


def main() -> int:
    objs = sorted((ROOT / "build" / "obj").glob("*.o"))
    if not objs:
        print("no objects under build/obj: run __graft_entry__.build() first", file=sys.stderr)
        return 1
    dest = ROOT / "profiles" / "sass"
    dest.mkdir(parents=True, exist_ok=True)
    for obj in objs:
        if obj.stem.startswith("common_") or obj.stem.startswith("engine_"):
            continue  # host-only translation units
        (dest / f"{obj.stem}.txt").write_text(summarize(obj))
        print(f"wrote profiles/sass/{obj.stem}.txt")
    return 0


if __name__ == "__main__":
    sys.exit(main())
