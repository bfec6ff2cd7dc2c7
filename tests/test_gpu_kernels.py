"""Memory-bound kernels + attention forward vs fp32 torch references (B200 box)."""

from __future__ import annotations

import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _bf(shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


def _rel(got, want):
    return ((got.float() - want.float()).abs().max() / (want.float().abs().max() + 1e-6)).item()


@pytest.mark.parametrize(
    "rows,d", [(64, 768), (1000, 1600), (37, 64), (4096, 1024), (2500, 2048), (3000, 8)]
)
def test_layernorm_fwd_bwd(rows: int, d: int) -> None:
    from llmtrain.ops import kernels as K

    x = _bf((rows, d), 2.0)
    gamma = torch.randn(d, device="cuda") * 0.5 + 1.0
    beta = torch.randn(d, device="cuda") * 0.1
    mean = torch.empty(rows, device="cuda")
    rstd = torch.empty(rows, device="cuda")
    y = K.layernorm_fwd(x, gamma, beta, mean=mean, rstd=rstd)

    xr = x.float().requires_grad_(True)
    g32, b32 = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    want = F.layer_norm(xr, (d,), g32, b32, 1e-5)
    assert _rel(y, want) < 1e-2
    assert _rel(mean, xr.mean(-1)) < 1e-3

    dy = _bf((rows, d))
    resid0 = _bf((rows, d))
    dresid = resid0.clone()
    dgamma = torch.zeros(d, device="cuda")
    dbeta = torch.zeros(d, device="cuda")
    K.layernorm_bwd(dy, x, gamma, mean, rstd, dresid, dgamma, dbeta)
    want.backward(dy.float())
    assert _rel(dresid, resid0.float() + xr.grad) < 2e-2
    assert _rel(dgamma, g32.grad) < 1e-2
    assert _rel(dbeta, b32.grad) < 1e-2
    # accumulation semantics of the parameter gradients
    K.layernorm_bwd(dy, x, gamma, mean, rstd, dresid, dgamma, dbeta)
    assert _rel(dgamma, 2 * g32.grad) < 1e-2
    # optional third output: column sums of the updated residual gradient
    dresid2 = resid0.clone()
    colsum = torch.ones(d, device="cuda")
    K.layernorm_bwd(dy, x, gamma, mean, rstd, dresid2, dgamma, dbeta, colsum)
    assert _rel(colsum, 1.0 + (resid0.float() + xr.grad).sum(0)) < 1e-2


def test_embedding_fwd_bwd() -> None:
    from llmtrain.ops import kernels as K

    b, t, d, v = 3, 50, 128, 1000
    ids = torch.randint(0, v, (b, t), device="cuda")
    wte, wpe = _bf((v, d)), _bf((64, d))
    x = K.embed_fwd(ids, wte, wpe)
    want = wte.float()[ids] + wpe.float()[:t][None]
    assert _rel(x.view(b, t, d), want) < 1e-2

    dx = _bf((b * t, d))
    dwte = torch.zeros(v, d, device="cuda")
    dwpe = torch.zeros(64, d, device="cuda")
    K.embed_bwd(ids, dx, dwte, dwpe)
    ref_wte = torch.zeros(v, d, device="cuda").index_add_(0, ids.flatten(), dx.float())
    assert _rel(dwte, ref_wte) < 1e-4
    assert _rel(dwpe[:t], dx.float().view(b, t, d).sum(0)) < 1e-4
    assert dwpe[t:].abs().max().item() == 0


@pytest.mark.parametrize(
    "rows,v,ldv",
    [(64, 1000, 1024), (33, 50257, 50304), (8, 16, 16), (1000, 5001, 5056), (5, 60000, 60032)],
)
def test_cross_entropy_loss_and_grad(rows: int, v: int, ldv: int) -> None:
    from llmtrain.ops import kernels as K

    logits = torch.zeros(rows, ldv, device="cuda", dtype=torch.bfloat16)
    logits[:, :v] = _bf((rows, v), 3.0)
    logits[:, v:] = 77.0  # padding columns must be ignored
    labels = torch.randint(0, v, (rows,), device="cuda")
    valid = torch.ones(rows, dtype=torch.uint8, device="cuda")
    valid[::5] = 0
    ref_in = logits[:, :v].float().requires_grad_(True)
    per_tok = F.cross_entropy(ref_in, labels, reduction="none")
    n_valid = int(valid.sum())
    (per_tok * valid.float()).sum().div(n_valid).backward()

    acc = torch.zeros(1, device="cuda")
    work = logits.clone()
    rows_loss = K.cross_entropy(work, labels, vocab=v, row_valid=valid, loss_acc=acc,
                                loss_scale=1.0 / n_valid, grad_scale=1.0 / n_valid)
    assert _rel(rows_loss, per_tok * valid.float()) < 1e-3
    assert abs(acc.item() - (per_tok * valid.float()).sum().item() / n_valid) < 1e-3
    assert _rel(work[:, :v], ref_in.grad) < 2e-2
    assert work[:, v:].abs().max().item() == 0 if ldv > v else True

    untouched = logits.clone()
    K.cross_entropy(untouched, labels, vocab=v, write_grad=False)
    assert torch.equal(untouched, logits)


@pytest.mark.parametrize("v,ldv", [(1000, 1024), (50257, 50304)])
def test_cross_entropy_ignore_index_rows_have_no_loss_and_no_gradient(v: int, ldv: int) -> None:
    """label == -100 follows ``F.cross_entropy``: zero loss, zero gradient, row still counted by
    the caller's mean (both kernels: the streaming one and the row-resident one)."""
    from llmtrain.ops import kernels as K

    rows = 48
    logits = torch.zeros(rows, ldv, device="cuda", dtype=torch.bfloat16)
    logits[:, :v] = _bf((rows, v), 2.0)
    labels = torch.randint(0, v, (rows,), device="cuda")
    labels[::3] = -100
    ref_in = logits[:, :v].float().requires_grad_(True)
    per_tok = F.cross_entropy(ref_in, labels, reduction="none")  # ignore_index=-100 -> 0
    (per_tok.sum() / rows).backward()
    work = logits.clone()
    rows_loss = K.cross_entropy(work, labels, vocab=v, grad_scale=1.0 / rows)
    assert _rel(rows_loss, per_tok) < 1e-3
    assert rows_loss[::3].abs().max().item() == 0
    assert work[::3].abs().max().item() == 0
    assert _rel(work[:, :v], ref_in.grad) < 2e-2


@pytest.mark.parametrize("m,n", [(4096, 768), (1000, 2304), (130, 50)])
def test_bias_grad(m: int, n: int) -> None:
    from llmtrain.ops import kernels as K

    dy = _bf((m, n))
    db = torch.ones(n, device="cuda")
    K.bias_grad(dy, db)
    assert _rel(db, 1.0 + dy.float().sum(0)) < 1e-3


def test_cast_and_sumsq_and_adamw_match_torch() -> None:
    from llmtrain.ops import kernels as K

    n = 1_000_003
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda") * 3
    shadow = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    K.cast_bf16(p, shadow)
    assert torch.equal(shadow, p.to(torch.bfloat16))

    sumsq = torch.zeros(1, device="cuda")
    K.grad_sumsq(g, sumsq)
    assert abs(sumsq.item() / g.double().pow(2).sum().item() - 1) < 1e-5

    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([ref_p], lr=1e-2, weight_decay=0.1)
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    ours = p.clone()
    norm = torch.zeros(1, device="cuda")
    for step in (1, 2, 3):
        g_step = g * step
        ref_p.grad = g_step.clone()
        ref_norm = torch.nn.utils.clip_grad_norm_([ref_p], 1.0)
        opt.step()
        work = g_step.clone()
        K.grad_sumsq(work, sumsq)
        K.adamw(ours, m, v, work, shadow, sumsq=sumsq, norm_out=norm, lr=1e-2, beta1=0.9,
                beta2=0.999, eps=1e-8, weight_decay=0.1, step=step, max_norm=1.0)
        assert abs(norm.item() / ref_norm.item() - 1) < 1e-4
        assert work.abs().max().item() == 0  # gradients zeroed for the next step
        assert (ours - ref_p.detach()).abs().max().item() < 1e-5
    assert torch.equal(shadow, ours.to(torch.bfloat16))
    assert _rel(m, opt.state[ref_p]["exp_avg"]) < 1e-5
    assert _rel(v, opt.state[ref_p]["exp_avg_sq"]) < 1e-5


def _ref_attention(qkv: torch.Tensor, h: int, key_mask: torch.Tensor | None):
    b, t, three_d = qkv.shape
    d = three_d // 3
    hd = d // h
    q, k, v = qkv.float().view(b, t, 3, h, hd).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
    lowest = torch.finfo(s.dtype).min
    s = s.masked_fill(torch.ones(t, t, dtype=torch.bool, device=qkv.device).triu_(1), lowest)
    if key_mask is not None:
        s = s.masked_fill(~key_mask.bool()[:, None, None, :], lowest)
    p = s.softmax(-1)
    return (p @ v).transpose(1, 2).reshape(b, t, d), torch.logsumexp(s, -1)


@pytest.mark.parametrize("b,t,h", [(2, 128, 2), (1, 256, 3), (2, 1024, 4), (3, 200, 2), (2, 8, 1)])
def test_attention_fwd(b: int, t: int, h: int) -> None:
    from llmtrain.ops import kernels as K

    qkv = _bf((b, t, 3 * h * 64))
    out, lse = K.attention_fwd(qkv, h)
    want, want_lse = _ref_attention(qkv, h, None)
    assert _rel(out, want) < 2e-2
    assert (lse - want_lse).abs().max().item() < 2e-2


def test_attention_fwd_key_padding() -> None:
    from llmtrain.ops import kernels as K

    b, t, h = 2, 300, 2
    qkv = _bf((b, t, 3 * h * 64))
    mask = torch.ones(b, t, dtype=torch.uint8, device="cuda")
    mask[0, 250:] = 0
    mask[1, 100:140] = 0
    out, _ = K.attention_fwd(qkv, h, key_mask=mask)
    want, _ = _ref_attention(qkv, h, mask)
    keep = mask.bool()[:, :, None]
    assert _rel(out * keep, want * keep) < 2e-2


@pytest.mark.parametrize("b,t,h", [(1, 128, 1), (2, 256, 2), (2, 1024, 3), (2, 200, 2), (1, 8, 2)])
def test_attention_bwd(b: int, t: int, h: int) -> None:
    from llmtrain.ops import kernels as K

    qkv = _bf((b, t, 3 * h * 64))
    out, lse = K.attention_fwd(qkv, h)
    dout = _bf((b, t, h * 64))
    dqkv = K.attention_bwd(qkv, out, dout, lse, h)

    ref_in = qkv.float().requires_grad_(True)
    want, _ = _ref_attention(ref_in, h, None)
    want.backward(dout.float())
    d = h * 64
    for name, sl in (("dq", slice(0, d)), ("dk", slice(d, 2 * d)), ("dv", slice(2 * d, 3 * d))):
        err = _rel(dqkv[..., sl], ref_in.grad[..., sl])
        assert err < 3e-2, f"{name}: {err}"


def test_attention_bwd_key_padding() -> None:
    from llmtrain.ops import kernels as K

    b, t, h = 2, 300, 2
    qkv = _bf((b, t, 3 * h * 64))
    mask = torch.ones(b, t, dtype=torch.uint8, device="cuda")
    mask[0, 250:] = 0
    mask[1, 100:140] = 0
    keep = mask.bool()[:, :, None]
    out, lse = K.attention_fwd(qkv, h, key_mask=mask)
    dout = (_bf((b, t, h * 64)).float() * keep).to(torch.bfloat16)  # padded queries get no grad
    dqkv = K.attention_bwd(qkv, out, dout, lse, h, key_mask=mask)
    ref_in = qkv.float().requires_grad_(True)
    want, _ = _ref_attention(ref_in, h, mask)
    want.backward(dout.float())
    assert _rel(dqkv, ref_in.grad) < 3e-2


@pytest.mark.parametrize("b,t,h", [(2, 384, 3), (1, 200, 2)])
def test_attention_bwd_accumulates_the_qkv_bias_gradient(b: int, t: int, h: int) -> None:
    """``dbias`` (what the engine passes) receives the column sums of the dqkv the same call
    writes, added to what it held (reference: ``qkv_proj``'s bias, `gpt.py:27`)."""
    from llmtrain.ops import kernels as K

    qkv = _bf((b, t, 3 * h * 64))
    out, lse = K.attention_fwd(qkv, h)
    dout = _bf((b, t, h * 64))
    plain = K.attention_bwd(qkv, out, dout, lse, h)
    dbias = torch.full((3 * h * 64,), 0.5, device="cuda")
    dqkv = K.attention_bwd(qkv, out, dout, lse, h, dbias=dbias)
    assert _rel(dqkv, plain) < 1e-3  # dq is summed by atomics: not bit-identical across calls
    want = dqkv.float().sum(dim=(0, 1)) + 0.5
    assert torch.allclose(dbias, want, rtol=1e-4, atol=1e-3), (dbias - want).abs().max()


@pytest.mark.parametrize("b,t,h", [(2, 384, 3), (2, 200, 2)])
def test_out_proj_dgrad_hands_the_row_statistics_to_attention_backward(b: int, t: int, h: int) -> None:
    """The engine's default path: out_proj's dgrad epilogue writes (LSE * log2 e, rowsum(dO o O));
    the attention backward then skips its own pass over dO and O and must give the same dqkv."""
    from llmtrain.ops import gemm as G
    from llmtrain.ops import kernels as K

    d = h * 64
    qkv = _bf((b, t, 3 * d))
    out, lse = K.attention_fwd(qkv, h)
    dy = _bf((b * t, d))
    w_out = (_bf((d, d)).float() / d ** 0.5).to(torch.bfloat16)
    d_out, dqkv, stats = K.attention_bwd_out_proj(dy, w_out, qkv, out, lse, h)

    want_dout = torch.empty(b * t, d, dtype=torch.bfloat16, device="cuda")
    G.gemm(G.NN, G.EPI_BIAS, dy, w_out, want_dout)
    assert torch.equal(d_out.view(b * t, d), want_dout)
    want_d = (d_out.float() * out.float()).view(b, t, h, 64).sum(-1).permute(0, 2, 1)  # [b, h, t]
    assert torch.allclose(stats[:, :, :t, 1], want_d, rtol=1e-4, atol=1e-3)
    assert torch.allclose(stats[:, :, :t, 0], lse * 1.4426950408889634, rtol=1e-6, atol=1e-6)
    if stats.shape[2] > t:  # padding rows: exact zeros for the probabilities
        assert torch.isinf(stats[:, :, t:, 0]).all() and (stats[:, :, t:, 1] == 0).all()
    want = K.attention_bwd(qkv, out, d_out, lse, h)
    assert _rel(dqkv, want) < 2e-3


def test_synthetic_tokens_are_shifted_in_range_and_reproducible() -> None:
    from llmtrain.ops import kernels as K

    ids, labels = K.synthetic_tokens(4, 257, 50257, seed=7, first_row=10)
    assert ids.shape == (4, 257) and ids.dtype == torch.int64
    assert int(ids.min()) >= 0 and int(ids.max()) < 50257
    assert torch.equal(labels[:, :-1], ids[:, 1:])
    again, _ = K.synthetic_tokens(4, 257, 50257, seed=7, first_row=10)
    assert torch.equal(ids, again)
    shifted, _ = K.synthetic_tokens(2, 257, 50257, seed=7, first_row=12)
    assert torch.equal(shifted, ids[2:])
    other, _ = K.synthetic_tokens(4, 257, 50257, seed=8, first_row=10)
    assert not torch.equal(ids, other)
    assert ids.float().mean().item() == pytest.approx(50257 / 2, rel=0.1)


def test_every_kernel_probe_runs_clean() -> None:
    """bench/sanitize_probe.py (the compute-sanitizer target) launches every native kernel once on
    tiny shapes and checks results; running it here keeps the target itself from rotting."""
    import importlib.util
    from pathlib import Path

    path = Path(__file__).resolve().parents[1] / "bench" / "sanitize_probe.py"
    spec = importlib.util.spec_from_file_location("sanitize_probe", path)
    assert spec is not None and spec.loader is not None
    module = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(module)
    assert module.main() == 0
