"""Wrappers for the non-GEMM sm_100a kernels (``include/llmtrain_b200.h``).

Each function launches on the current CUDA stream and writes into caller-provided tensors (the
C++ engine does the same with arena pointers); the small ``scratch`` buffers needed by the
deterministic two-stage reductions are cached per device.
"""

from __future__ import annotations

import ctypes
import math

import torch

from llmtrain.ops import native
from llmtrain.ops.native import check, ptr, stream_handle

_c_int = ctypes.c_int
_c_float = ctypes.c_float
_c_size = ctypes.c_size_t

_scratch: dict[int, torch.Tensor] = {}


def scratch(nbytes: int, device: torch.device) -> torch.Tensor:
    """Per-device byte buffer of at least ``nbytes`` (grown geometrically, never shrunk)."""
    index = device.index if device.index is not None else torch.cuda.current_device()
    buf = _scratch.get(index)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 22), dtype=torch.uint8, device=device)
        _scratch[index] = buf
    return buf


def _lib():
    lib = native.lib()
    if not getattr(lib, "_llm_typed", False):
        for name in (
            "llmtrain_b200_layernorm_bwd_scratch_bytes",
            "llmtrain_b200_bias_grad_scratch_bytes",
            "llmtrain_b200_grad_sumsq_scratch_bytes",
            "llmtrain_b200_attention_bwd_scratch_bytes",
        ):
            getattr(lib, name).restype = ctypes.c_size_t
        lib._llm_typed = True
    return lib


def layernorm_fwd(x, gamma, beta, y=None, mean=None, rstd=None):
    rows, d = x.shape
    y = torch.empty_like(x) if y is None else y
    check(
        _lib().llmtrain_b200_layernorm_fwd(
            ptr(x), ptr(gamma), ptr(beta), ptr(y), ptr(mean), ptr(rstd), _c_int(rows), _c_int(d),
            stream_handle(),
        ),
        "layernorm_fwd",
    )
    return y


def layernorm_bwd(dy, x, gamma, mean, rstd, dresid, dgamma, dbeta, dcolsum=None):
    rows, d = x.shape
    lib = _lib()
    need = lib.llmtrain_b200_layernorm_bwd_scratch_bytes(_c_int(rows), _c_int(d))
    buf = scratch(need, x.device)
    check(
        lib.llmtrain_b200_layernorm_bwd(
            ptr(dy), ptr(x), ptr(gamma), ptr(mean), ptr(rstd), ptr(dresid), ptr(dgamma),
            ptr(dbeta), ptr(dcolsum), ptr(buf), _c_size(buf.numel()), _c_int(rows), _c_int(d),
            stream_handle(),
        ),
        "layernorm_bwd",
    )


def embed_fwd(ids, wte, wpe, out=None):
    b, t = ids.shape
    vocab, d = wte.shape
    out = torch.empty(b * t, d, dtype=torch.bfloat16, device=ids.device) if out is None else out
    check(
        _lib().llmtrain_b200_embed_fwd(
            ptr(ids), ptr(wte), ptr(wpe), ptr(out), _c_int(b), _c_int(t), _c_int(d), _c_int(vocab),
            stream_handle(),
        ),
        "embed_fwd",
    )
    return out


def embed_bwd(ids, dx, dwte, dwpe):
    b, t = ids.shape
    vocab, d = dwte.shape
    check(
        _lib().llmtrain_b200_embed_bwd(
            ptr(ids), ptr(dx), ptr(dwte), ptr(dwpe), _c_int(b), _c_int(t), _c_int(d), _c_int(vocab),
            stream_handle(),
        ),
        "embed_bwd",
    )


def cross_entropy(logits, labels, *, vocab, row_valid=None, loss_rows=None, loss_acc=None,
                  loss_scale=1.0, grad_scale=1.0, write_grad=True):
    rows, ldv = logits.shape[0], logits.stride(0)
    if loss_rows is None:
        loss_rows = torch.empty(rows, dtype=torch.float32, device=logits.device)
    check(
        _lib().llmtrain_b200_cross_entropy(
            ptr(logits), ptr(labels), ptr(row_valid), ptr(loss_rows), ptr(loss_acc), _c_int(rows),
            _c_int(vocab), _c_int(ldv), _c_float(loss_scale), _c_float(grad_scale),
            _c_int(1 if write_grad else 0), stream_handle(),
        ),
        "cross_entropy",
    )
    return loss_rows


def bias_grad(dy, dbias):
    m, n = dy.shape
    lib = _lib()
    need = lib.llmtrain_b200_bias_grad_scratch_bytes(_c_int(m), _c_int(n))
    buf = scratch(need, dy.device)
    check(
        lib.llmtrain_b200_bias_grad(
            ptr(dy), _c_int(dy.stride(0)), _c_int(m), _c_int(n), ptr(dbias), ptr(buf),
            _c_size(buf.numel()), stream_handle(),
        ),
        "bias_grad",
    )


def cast_bf16(src, dst):
    check(
        _lib().llmtrain_b200_cast_f32_to_bf16(ptr(src), ptr(dst), _c_size(src.numel()),
                                              stream_handle()),
        "cast_f32_to_bf16",
    )


def grad_sumsq(g, out, *, accumulate=False):
    lib = _lib()
    need = lib.llmtrain_b200_grad_sumsq_scratch_bytes(_c_size(g.numel()))
    buf = scratch(need, g.device)
    check(
        lib.llmtrain_b200_grad_sumsq(
            ptr(g), _c_size(g.numel()), ptr(out), _c_int(1 if accumulate else 0), ptr(buf),
            _c_size(buf.numel()), stream_handle(),
        ),
        "grad_sumsq",
    )


def adamw(p, m, v, g, p16, *, sumsq, norm_out, lr, beta1, beta2, eps, weight_decay, step,
          grad_scale=1.0, max_norm=0.0, zero_grad=True):
    check(
        _lib().llmtrain_b200_adamw(
            ptr(p), ptr(m), ptr(v), ptr(g), ptr(p16), _c_size(p.numel()), ptr(sumsq),
            ptr(norm_out), _c_float(lr), _c_float(beta1), _c_float(beta2), _c_float(eps),
            _c_float(weight_decay), _c_int(step), _c_float(grad_scale), _c_float(max_norm),
            _c_int(1 if zero_grad else 0), stream_handle(),
        ),
        "adamw",
    )


def _c_u64(v: int) -> ctypes.c_uint64:
    return ctypes.c_uint64(int(v) % (1 << 64))


def dropout(x, p, *, seed, offset, site, out=None):
    """``out = x * mask / (1 - p)`` on a bf16 ``(rows, cols)`` tensor (counter-based mask)."""
    out = torch.empty_like(x) if out is None else out
    rows, cols = x.shape
    check(
        _lib().llmtrain_b200_dropout(
            ptr(x), ptr(out), ctypes.c_longlong(rows), _c_int(cols), _c_int(x.stride(0)),
            _c_float(p), _c_u64(seed), _c_u64(offset), ctypes.c_uint32(site), stream_handle(),
        ),
        "dropout",
    )
    return out


def dropout_mask(rows, cols, p, *, seed, offset, site, device="cuda"):
    """uint8 ``(rows, cols)`` keep-mask the fused kernels apply for these counters."""
    out = torch.empty(rows, cols, dtype=torch.uint8, device=device)
    check(
        _lib().llmtrain_b200_dropout_mask(
            ptr(out), ctypes.c_longlong(rows), _c_int(cols), _c_float(p), _c_u64(seed),
            _c_u64(offset), ctypes.c_uint32(site), stream_handle(),
        ),
        "dropout_mask",
    )
    return out


def dropout_scale(p: float) -> float:
    """The exact factor kept elements are multiplied by (p is quantised to 1/65536)."""
    if p <= 0:
        return 1.0
    t = min(max(int(p * 65536.0 + 0.5), 1), 65535)
    return 65536.0 / (65536 - t)


def linear_dropout_resid(x, w, bias, resid, p, *, seed, offset, site, out=None):
    """``dropout(x @ w.T + bias) + resid`` — the residual-branch projections with dropout."""
    m, k = x.shape
    n = w.shape[0]
    out = torch.empty(m, n, dtype=torch.bfloat16, device=x.device) if out is None else out
    check(
        _lib().llmtrain_b200_gemm_dropout_resid(
            _c_int(m), _c_int(n), _c_int(k), ptr(x), _c_int(x.stride(0)), ptr(w),
            _c_int(w.stride(0)), ptr(out), _c_int(out.stride(0)), ptr(bias), ptr(resid),
            _c_int(resid.stride(0)), _c_float(p), _c_u64(seed), _c_u64(offset),
            ctypes.c_uint32(site), stream_handle(),
        ),
        "gemm_dropout_resid",
    )
    return out


def attention_fwd(qkv, n_heads, *, key_mask=None, out=None, lse=None, dropout=None):
    """``qkv``: (B, T, 3*d) bf16 → ``(out (B, T, d) bf16, lse (B, H, T) fp32)``.

    ``dropout``: optional ``(p, seed, offset, site)`` for dropout on the attention probabilities.
    """
    b, t, three_d = qkv.shape
    d = three_d // 3
    hd = d // n_heads
    out = torch.empty(b, t, d, dtype=torch.bfloat16, device=qkv.device) if out is None else out
    if lse is None:
        lse = torch.empty(b, n_heads, t, dtype=torch.float32, device=qkv.device)
    if dropout is not None:
        p, seed, offset, site = dropout
        check(
            _lib().llmtrain_b200_attention_fwd_dropout(
                ptr(qkv), ptr(out), ptr(lse), ptr(key_mask), _c_int(b), _c_int(t),
                _c_int(n_heads), _c_int(hd), _c_float(p), _c_u64(seed), _c_u64(offset),
                ctypes.c_uint32(site), stream_handle(),
            ),
            "attention_fwd_dropout",
        )
        return out, lse
    check(
        _lib().llmtrain_b200_attention_fwd(
            ptr(qkv), ptr(out), ptr(lse), ptr(key_mask), _c_int(b), _c_int(t), _c_int(n_heads),
            _c_int(hd), stream_handle(),
        ),
        "attention_fwd",
    )
    return out, lse


def attention_bwd(qkv, out, dout, lse, n_heads, *, key_mask=None, dqkv=None, dropout=None,
                  dbias=None):
    """Gradient of :func:`attention_fwd` w.r.t. the fused ``qkv`` buffer.

    ``dbias`` (fp32 ``[3 * d_model]``, optional): the column sums of ``dqkv`` — ``qkv_proj``'s bias
    gradient — are ADDED to it by the same kernels (what the training engine does).
    """
    b, t, three_d = qkv.shape
    hd = three_d // 3 // n_heads
    dqkv = torch.empty_like(qkv) if dqkv is None else dqkv
    lib = _lib()
    need = lib.llmtrain_b200_attention_bwd_scratch_bytes(_c_int(b), _c_int(t), _c_int(n_heads),
                                                         _c_int(hd))
    buf = scratch(need, qkv.device)
    if dbias is not None:
        assert dbias.dtype == torch.float32 and dbias.numel() == three_d and dbias.is_contiguous()
        p, seed, offset, site = dropout if dropout is not None else (0.0, 0, 0, 0)
        check(
            lib.llmtrain_b200_attention_bwd_fused(
                ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(key_mask), ptr(dqkv), ptr(dbias),
                ptr(buf), _c_size(buf.numel()), _c_int(b), _c_int(t), _c_int(n_heads), _c_int(hd),
                _c_float(0.0), _c_float(p), _c_u64(seed), _c_u64(offset), ctypes.c_uint32(site),
                stream_handle(),
            ),
            "attention_bwd_fused",
        )
        return dqkv
    if dropout is not None:
        p, seed, offset, site = dropout
        check(
            lib.llmtrain_b200_attention_bwd_dropout(
                ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(key_mask), ptr(dqkv), ptr(buf),
                _c_size(buf.numel()), _c_int(b), _c_int(t), _c_int(n_heads), _c_int(hd),
                _c_float(p), _c_u64(seed), _c_u64(offset), ctypes.c_uint32(site), stream_handle(),
            ),
            "attention_bwd_dropout",
        )
        return dqkv
    check(
        lib.llmtrain_b200_attention_bwd(
            ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(key_mask), ptr(dqkv), ptr(buf),
            _c_size(buf.numel()), _c_int(b), _c_int(t), _c_int(n_heads), _c_int(hd),
            stream_handle(),
        ),
        "attention_bwd",
    )
    return dqkv


def attention_bwd_out_proj(dy, w_out, qkv, out, lse, n_heads, *, key_mask=None, dbias=None):
    """out_proj's dgrad fused with the attention backward's statistics pre-pass, then the attention
    backward itself (what the training engine runs without probability dropout).

    ``dy`` ``[B*T, d_model]`` is the gradient w.r.t. out_proj's output, ``w_out`` ``[d_model, H*64]``
    its weight.  Returns ``(d_out [B, T, H*64], dqkv, stats [B, H, t_stat, 2])``.
    """
    from llmtrain.ops import gemm as G

    b, t, three_d = qkv.shape
    hd = three_d // 3 // n_heads
    lib = _lib()
    need = lib.llmtrain_b200_attention_bwd_scratch_bytes(_c_int(b), _c_int(t), _c_int(n_heads),
                                                         _c_int(hd))
    buf = scratch(need, qkv.device)
    stats_ptr, t_stat = ctypes.c_void_p(), ctypes.c_int()
    check(lib.llmtrain_b200_attention_bwd_stats_buffer(
        ptr(buf), _c_size(buf.numel()), _c_int(b), _c_int(t), _c_int(n_heads), _c_int(hd),
        ctypes.byref(stats_ptr), ctypes.byref(t_stat)), "attention_bwd_stats_buffer")
    offset = stats_ptr.value - buf.data_ptr()
    stats = buf[offset:offset + b * n_heads * t_stat.value * 8].view(torch.float32).view(
        b, n_heads, t_stat.value, 2)
    d_out = torch.empty(b * t, n_heads * hd, dtype=torch.bfloat16, device=qkv.device)
    G.gemm_rowdot(dy, w_out, d_out, out.view(b * t, -1), lse, stats, t, n_heads)
    dqkv = torch.empty_like(qkv)
    check(
        lib.llmtrain_b200_attention_bwd_prestats(
            ptr(qkv), ptr(out), ptr(d_out), ptr(lse), ptr(key_mask), ptr(dqkv), ptr(dbias),
            ptr(buf), _c_size(buf.numel()), _c_int(b), _c_int(t), _c_int(n_heads), _c_int(hd),
            _c_float(0.0), stream_handle(),
        ),
        "attention_bwd_prestats",
    )
    return d_out.view(b, t, -1), dqkv, stats


def synthetic_tokens(batch: int, seq: int, vocab: int, *, seed: int, first_row: int = 0,
                     device: torch.device | str = "cuda", with_labels: bool = True):
    """Device-resident ``(input_ids, labels)`` (next-token targets) from a counter-based hash."""
    ids = torch.empty(batch, seq, dtype=torch.int64, device=device)
    labels = torch.empty_like(ids) if with_labels else None
    check(
        _lib().llmtrain_b200_synthetic_tokens(
            ptr(ids), ptr(labels), _c_int(batch), _c_int(seq), _c_int(vocab),
            ctypes.c_uint64(seed), ctypes.c_uint64(first_row), stream_handle()),
        "synthetic_tokens",
    )
    return ids, labels


def softmax_scale(head_dim: int) -> float:
    return 1.0 / math.sqrt(head_dim)
