"""tcgen05 GEMM family (``csrc/kernels/gemm.cu``) — see ``llmtrain_b200_gemm`` in the header."""

from __future__ import annotations

import ctypes

import torch

from llmtrain.ops import native

NT, NN, TN = 0, 1, 2
EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_RESID, EPI_DGELU, EPI_ACCUM_F32 = 0, 1, 2, 3, 4


def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, "operands must be 2-D with unit inner stride"
    return int(t.stride(0))


def gemm(
    layout: int,
    epilogue: int,
    a: torch.Tensor,
    b: torch.Tensor,
    c: torch.Tensor,
    *,
    c2: torch.Tensor | None = None,
    bias: torch.Tensor | None = None,
    r: torch.Tensor | None = None,
) -> torch.Tensor:
    """Run one GEMM on the current stream; ``c`` (and ``c2``) are written in place."""
    m, n = c.shape
    k = a.shape[1] if layout != TN else a.shape[0]
    fn = native.lib().llmtrain_b200_gemm
    status = fn(
        ctypes.c_int(layout), ctypes.c_int(epilogue), ctypes.c_int(m), ctypes.c_int(n),
        ctypes.c_int(k), native.ptr(a), ctypes.c_int(_ld(a)), native.ptr(b), ctypes.c_int(_ld(b)),
        native.ptr(c), ctypes.c_int(_ld(c)), native.ptr(c2),
        ctypes.c_int(_ld(c2) if c2 is not None else 0), native.ptr(bias), native.ptr(r),
        ctypes.c_int(_ld(r) if r is not None else 0), native.stream_handle(),
    )
    native.check(status, "gemm")
    return c


def gemm_colsum(
    layout: int,
    epilogue: int,
    a: torch.Tensor,
    b: torch.Tensor,
    c: torch.Tensor,
    colsum: torch.Tensor,
    *,
    bias: torch.Tensor | None = None,
    r: torch.Tensor | None = None,
) -> torch.Tensor:
    """As :func:`gemm` (plain or dGELU epilogue) and ``colsum += c.sum(0)`` from the same epilogue."""
    m, n = c.shape
    k = a.shape[1] if layout != TN else a.shape[0]
    status = native.lib().llmtrain_b200_gemm_colsum(
        ctypes.c_int(layout), ctypes.c_int(epilogue), ctypes.c_int(m), ctypes.c_int(n),
        ctypes.c_int(k), native.ptr(a), ctypes.c_int(_ld(a)), native.ptr(b), ctypes.c_int(_ld(b)),
        native.ptr(c), ctypes.c_int(_ld(c)), native.ptr(bias), native.ptr(r),
        ctypes.c_int(_ld(r) if r is not None else 0), native.ptr(colsum), native.stream_handle(),
    )
    native.check(status, "gemm_colsum")
    return c


def gemm_rowdot(dy: torch.Tensor, w: torch.Tensor, d_out: torch.Tensor, o: torch.Tensor,
                lse: torch.Tensor, stats: torch.Tensor, seq_len: int, n_heads: int) -> torch.Tensor:
    """``d_out[M, H*64] = dy[M, K] @ w[K, H*64]`` (out_proj's dgrad) and, from the same epilogue,
    ``stats[b, h, t] = (lse[b, h, t] * log2 e, sum_c d_out[b, t, h, c] * o[b, t, h, c])``.

    ``stats``: fp32 ``[B, H, t_stat, 2]`` with ``t_stat >= seq_len`` (entries ``t >= seq_len`` are
    left alone); rows of ``dy`` / ``o`` are ``(b, t)``.
    """
    m, n = d_out.shape
    k = dy.shape[1]
    assert stats.dtype == torch.float32 and stats.shape[-1] == 2 and stats.is_contiguous()
    status = native.lib().llmtrain_b200_gemm_rowdot(
        ctypes.c_int(m), ctypes.c_int(n), ctypes.c_int(k), native.ptr(dy), ctypes.c_int(_ld(dy)),
        native.ptr(w), ctypes.c_int(_ld(w)), native.ptr(d_out), ctypes.c_int(_ld(d_out)),
        native.ptr(o), ctypes.c_int(_ld(o)), native.ptr(lse), native.ptr(stats),
        ctypes.c_int(seq_len), ctypes.c_int(n_heads), ctypes.c_int(int(stats.shape[-2])),
        native.stream_handle(),
    )
    native.check(status, "gemm_rowdot")
    return d_out


def linear_forward(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """``x @ w.T + bias`` with bf16 operands, fp32 accumulate, bf16 result."""
    out = torch.empty(x.shape[0], w.shape[0], dtype=torch.bfloat16, device=x.device)
    return gemm(NT, EPI_BIAS, x, w, out, bias=bias)
