"""ctypes binding to ``libllmtrain_b200.so`` (C ABI: ``include/llmtrain_b200.h``).

Loading is lazy and loud: on a machine with a CUDA device a missing or unloadable library is an
error (never a silent fall-back to eager torch ops); on a CPU-only machine importing this module
is harmless and :func:`lib` raises only when a native op is actually requested.
"""

from __future__ import annotations

import ctypes
from pathlib import Path
from typing import Any

import torch  # noqa: F401  (maps libcudart before our library asks for it)

import os

# LLMTRAIN_NATIVE_LIB points at an alternative build of the library (A/B measurements of two
# kernel versions inside one process tree / one GPU box); the default is the in-tree build.
_LIB_PATH = Path(
    os.environ.get("LLMTRAIN_NATIVE_LIB")
    or Path(__file__).resolve().parents[1] / "_native" / "libllmtrain_b200.so"
)
_lib: ctypes.CDLL | None = None


class NativeError(RuntimeError):
    """A native entry point returned non-zero."""


def library_path() -> Path:
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    """The loaded library (built on demand if the ``.so`` is missing and nvcc is around)."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        from llmtrain.ops.build import build

        build()
    handle = ctypes.CDLL(str(_LIB_PATH), mode=ctypes.RTLD_GLOBAL)
    handle.llmtrain_b200_last_error.restype = ctypes.c_char_p
    handle.llmtrain_b200_abi_version.restype = ctypes.c_int
    _lib = handle
    return handle


def check(status: int, what: str) -> None:
    if status != 0:
        detail = lib().llmtrain_b200_last_error()
        raise NativeError(f"{what} failed: {detail.decode() if detail else 'unknown error'}")


def ptr(t: Any) -> ctypes.c_void_p:
    """Device pointer of a tensor (``None`` -> NULL)."""
    if t is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(t.data_ptr())


def stream_handle() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
