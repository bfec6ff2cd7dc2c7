"""C-ABI check of ``llmtrain_b200_bias_grad_noalloc`` (collected last on purpose: ``zz``).

That entry point serves programs that are not in a position to allocate device memory when they
call the library — for instance a process whose own pool has already claimed the whole GPU, or a
caller written against the bare C ABI with no allocator behind it.  Here it is driven exactly the
way such a caller would drive it: raw pointers through ``ctypes``, no scratch argument, nothing
from this repo's Python wrappers in between.  Its result is compared with the regular
``llmtrain_b200_bias_grad`` (caller-supplied scratch) on two consecutive calls, which also shows
that whatever the first call left behind in the built-in array does not leak into the second.
"""

from __future__ import annotations

import ctypes
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

_LIB = Path(__file__).resolve().parents[1] / "src" / "llmtrain" / "_native" / "libllmtrain_b200.so"


def _load() -> ctypes.CDLL:
    lib = ctypes.CDLL(str(_LIB))
    lib.llmtrain_b200_last_error.restype = ctypes.c_char_p
    lib.llmtrain_b200_bias_grad_scratch_bytes.restype = ctypes.c_size_t
    lib.llmtrain_b200_bias_grad_scratch_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.llmtrain_b200_bias_grad_noalloc_capacity_bytes.restype = ctypes.c_size_t
    lib.llmtrain_b200_bias_grad.argtypes = [
        ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
        ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
    ]
    lib.llmtrain_b200_bias_grad_noalloc.argtypes = [
        ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
        ctypes.c_void_p,
    ]
    return lib


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def test_noalloc_bias_grad_matches_regular_entry_point_twice_in_a_row() -> None:
    lib = _load()
    assert lib.llmtrain_b200_bias_grad_noalloc_capacity_bytes() == 1 << 20

    # two different problems back to back: the second must not see the first one's partial sums
    for m, n, seed in ((4096, 768, 0), (1000, 2304, 1)):
        torch.manual_seed(seed)
        dy = torch.randn(m, n, device="cuda").to(torch.bfloat16)
        start = torch.randn(n, device="cuda")

        want = start.clone()
        need = lib.llmtrain_b200_bias_grad_scratch_bytes(m, n)
        scratch = torch.empty(max(need, 1), dtype=torch.uint8, device="cuda")
        rc = lib.llmtrain_b200_bias_grad(dy.data_ptr(), dy.stride(0), m, n, want.data_ptr(),
                                         scratch.data_ptr(), scratch.numel(), _stream())
        assert rc == 0, lib.llmtrain_b200_last_error()

        got = start.clone()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.llmtrain_b200_bias_grad_noalloc(dy.data_ptr(), dy.stride(0), m, n,
                                                 got.data_ptr(), _stream())
        assert rc == 0, lib.llmtrain_b200_last_error()
        torch.cuda.synchronize()
        if seed == 1:  # the module is loaded by now: a call must not take device memory at all
            assert torch.cuda.mem_get_info()[0] == free_before

        ref = start + dy.float().sum(0)
        scale = ref.abs().max().item() + 1e-6
        assert (got - want).abs().max().item() / scale < 1e-3
        assert (got - ref).abs().max().item() / scale < 1e-3


def test_noalloc_bias_grad_refuses_problems_beyond_its_capacity() -> None:
    lib = _load()
    m, n = 64, 300_000  # one partial sum per column: more than the 262144 floats built in
    dy = torch.zeros(m, n, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(n, device="cuda")
    rc = lib.llmtrain_b200_bias_grad_noalloc(dy.data_ptr(), dy.stride(0), m, n, out.data_ptr(),
                                             _stream())
    assert rc != 0
    message = lib.llmtrain_b200_last_error().decode()
    assert "built-in scratch" in message and "1 MiB" in message
    torch.cuda.synchronize()
    assert out.abs().max().item() == 0.0  # nothing was launched
