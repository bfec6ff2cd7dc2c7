"""Build ``libllmtrain_b200.so`` in-tree with nvcc for sm_100a (cross-compiles without a GPU).

The library is plain CUDA C++ with a C ABI (``include/llmtrain_b200.h``): no torch headers, so a
full rebuild takes seconds, and the resulting ``.so`` travels with the source tree.
"""

from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PACKAGE_DIR = Path(__file__).resolve().parents[1]
REPO_ROOT = PACKAGE_DIR.parents[1]
CSRC = REPO_ROOT / "csrc"
INCLUDE = REPO_ROOT / "include"
BUILD_DIR = REPO_ROOT / "build" / "obj"
LIB_PATH = PACKAGE_DIR / "_native" / "libllmtrain_b200.so"

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-O3",
    "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler",
    "-fPIC",
]


def _nvcc() -> str:
    found = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(found).exists():
        raise RuntimeError("nvcc not found; cannot build the native library")
    return found


def sources() -> list[Path]:
    return sorted(CSRC.rglob("*.cu"))


def _digest(path: Path, headers: list[Path]) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for item in [path, *headers]:
        h.update(item.read_bytes())
    return h.hexdigest()[:16]


def _compile_one(src: Path, headers: list[Path], verbose: bool) -> Path:
    rel = src.relative_to(CSRC).with_suffix("")
    obj = BUILD_DIR / ("_".join(rel.parts) + ".o")
    stamp = obj.with_suffix(".sha")
    digest = _digest(src, headers)
    if obj.exists() and stamp.exists() and stamp.read_text() == digest:
        return obj
    cmd = [_nvcc(), *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        sys.stderr.write(proc.stderr)
    stamp.write_text(digest)
    return obj


def build(*, verbose: bool = False, force: bool = False) -> Path:
    """Compile every ``csrc/**/*.cu`` for sm_100a and link the shared library."""
    BUILD_DIR.mkdir(parents=True, exist_ok=True)
    LIB_PATH.parent.mkdir(parents=True, exist_ok=True)
    if force:
        for stale in BUILD_DIR.glob("*.sha"):
            stale.unlink()
    headers = sorted([*CSRC.rglob("*.cuh"), *CSRC.rglob("*.h"), *INCLUDE.glob("*.h")])
    srcs = sources()
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as pool:
        objs = list(pool.map(lambda s: _compile_one(s, headers, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if LIB_PATH.exists() and LIB_PATH.stat().st_mtime >= newest and not force:
        return LIB_PATH
    link = [
        _nvcc(),
        "-shared",
        "-gencode",
        "arch=compute_100a,code=sm_100a",
        "-Xcompiler",
        "-fPIC",
        "-o",
        str(LIB_PATH),
        *map(str, objs),
        # Share torch's CUDA runtime instance (same current-device state, same stream handles):
        # libcudart.so.12 is already mapped once torch is imported; the rpath is the fallback.
        "--cudart",
        "shared",
        "-Xlinker",
        "-rpath=/usr/local/cuda/lib64",
    ]
    proc = subprocess.run(link, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"link failed:\n{proc.stdout}\n{proc.stderr}")
    return LIB_PATH


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="--force" in sys.argv))
