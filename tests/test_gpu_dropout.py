"""Native dropout (counter-based masks) vs PyTorch fp32 references that apply the SAME mask.

The reference keeps four nn.Dropout modules (`src/llmtrain/models/gpt.py:28,30,97,129`); here the
mask is a function of (seed, offset, site, element index), dumped with `dropout_mask` so the torch
side of each comparison can apply it explicitly.
"""

from __future__ import annotations

import math

import pytest
import torch

pytestmark = pytest.mark.gpu

SEED, OFFSET = 0x1234_5678_9ABC_DEF0, 41


def _bf(shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


def _rel(a: torch.Tensor, b: torch.Tensor) -> float:
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


@pytest.mark.parametrize("p", [0.1, 0.5])
def test_mask_statistics_and_determinism(p: float) -> None:
    from llmtrain.ops import kernels as K

    m1 = K.dropout_mask(4096, 768, p, seed=SEED, offset=OFFSET, site=5)
    m2 = K.dropout_mask(4096, 768, p, seed=SEED, offset=OFFSET, site=5)
    assert torch.equal(m1, m2)
    keep = m1.float().mean().item()
    assert abs(keep - (1 - p)) < 3e-3
    # different site / offset / seed -> (practically) independent masks
    for other in (
        K.dropout_mask(4096, 768, p, seed=SEED, offset=OFFSET, site=6),
        K.dropout_mask(4096, 768, p, seed=SEED, offset=OFFSET + 1, site=5),
        K.dropout_mask(4096, 768, p, seed=SEED + 1, offset=OFFSET, site=5),
    ):
        agree = (other == m1).float().mean().item()
        expect = (1 - p) ** 2 + p**2
        assert abs(agree - expect) < 5e-3
    # no structure along rows or columns
    assert (m1.float().mean(0) - (1 - p)).abs().max().item() < 0.05
    assert (m1.float().mean(1) - (1 - p)).abs().max().item() < 0.08


def test_elementwise_dropout_matches_mask_and_handles_ragged_width() -> None:
    from llmtrain.ops import kernels as K

    for rows, cols, p in ((513, 768, 0.2), (37, 50, 0.3)):
        x = _bf((rows, cols))
        y = K.dropout(x, p, seed=SEED, offset=OFFSET, site=2)
        mask = K.dropout_mask(rows, cols, p, seed=SEED, offset=OFFSET, site=2).float()
        want = x.float() * mask * K.dropout_scale(p)
        assert _rel(y, want) < 1e-2
        assert torch.equal(y == 0, (want == 0))
    # p = 0 is the identity
    x = _bf((64, 128))
    assert torch.equal(K.dropout(x, 0.0, seed=1, offset=2, site=3), x)


@pytest.mark.parametrize("m,n,k", [(512, 768, 768), (300, 768, 3072), (1024, 256, 128)])
def test_gemm_dropout_resid_epilogue(m: int, n: int, k: int) -> None:
    from llmtrain.ops import kernels as K

    p = 0.25
    x, w = _bf((m, k)), _bf((n, k), 0.05)
    bias = torch.randn(n, device="cuda")
    resid = _bf((m, n))
    out = K.linear_dropout_resid(x, w, bias, resid, p, seed=SEED, offset=OFFSET, site=9)
    mask = K.dropout_mask(m, n, p, seed=SEED, offset=OFFSET, site=9).float()
    want = (x.float() @ w.float().T + bias) * mask * K.dropout_scale(p) + resid.float()
    assert _rel(out, want) < 2e-2
    # dropped positions carry exactly the residual
    dropped = mask == 0
    assert torch.equal(out[dropped], resid[dropped])


def _ref_attention_dropout(qkv, h, keep_mask, scale):
    b, t, three_d = qkv.shape
    d = three_d // 3
    hd = d // h
    q, k, v = qkv.view(b, t, 3, h, hd).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
    s = s.masked_fill(torch.ones(t, t, dtype=torch.bool, device=qkv.device).triu_(1),
                      torch.finfo(s.dtype).min)
    probs = s.softmax(-1) * keep_mask.view(b, h, t, t) * scale
    return (probs @ v).transpose(1, 2).reshape(b, t, d)


@pytest.mark.parametrize("b,t,h", [(2, 256, 2), (1, 1024, 2), (2, 200, 1)])
def test_attention_dropout_fwd_bwd(b: int, t: int, h: int) -> None:
    from llmtrain.ops import kernels as K

    p, site = 0.2, 4
    drop = (p, SEED, OFFSET, site)
    qkv = _bf((b, t, 3 * h * 64))
    out, lse = K.attention_fwd(qkv, h, dropout=drop)
    keep = K.dropout_mask(b * h * t, t, p, seed=SEED, offset=OFFSET, site=site).float()
    ref_in = qkv.float().requires_grad_(True)
    want = _ref_attention_dropout(ref_in, h, keep, K.dropout_scale(p))
    assert _rel(out, want) < 2e-2
    # the saved log-sum-exp is that of the UNdropped softmax
    _, lse_plain = K.attention_fwd(qkv, h)
    assert (lse - lse_plain).abs().max().item() < 1e-4

    dout = _bf((b, t, h * 64))
    dqkv = K.attention_bwd(qkv, out, dout, lse, h, dropout=drop)
    want.backward(dout.float())
    d = h * 64
    for name, sl in (("dq", slice(0, d)), ("dk", slice(d, 2 * d)), ("dv", slice(2 * d, 3 * d))):
        err = _rel(dqkv[..., sl], ref_in.grad[..., sl])
        assert err < 3e-2, f"{name}: {err}"


def test_engine_dropout_training_is_reproducible_and_off_in_eval() -> None:
    """Whole model: train-mode losses depend on the mask counters, eval ignores dropout."""
    from llmtrain.models.gpt import GPT

    def build() -> GPT:
        torch.manual_seed(7)
        return GPT(vocab_size=512, block_size=128, d_model=128, n_layers=2, n_heads=2, d_ff=512,
                   dropout=0.3, tie_embeddings=True).cuda()

    ids = torch.randint(0, 512, (4, 128), device="cuda")
    labels = torch.randint(0, 512, (4, 128), device="cuda")

    def losses_and_grad(model: GPT, n: int):
        model.train()
        eng = model.native_engine()
        out = []
        for _ in range(n):
            model.zero_grad(set_to_none=True)
            loss = eng.loss(ids, labels, None)
            loss.backward()
            out.append(loss.item())
        g = model.token_embedding.weight.grad.detach().clone()
        return out, g

    a, ga = losses_and_grad(build(), 3)
    b, gb = losses_and_grad(build(), 3)
    # same seed, same counters -> same masks (gradients only differ by fp32 reduction order)
    assert a == b and _rel(ga, gb) < 1e-3
    assert len(set(a)) == 3  # a new mask every call, so the loss moves
    assert all(math.isfinite(v) for v in a) and torch.isfinite(ga).all()

    model = build()
    model.eval()
    eng = model.native_engine()
    with torch.no_grad():
        e1 = eng.loss(ids, labels, None).item()
        e2 = eng.loss(ids, labels, None).item()
    assert e1 == e2
    plain = GPT(vocab_size=512, block_size=128, d_model=128, n_layers=2, n_heads=2, d_ff=512,
                dropout=0.0, tie_embeddings=True).cuda()
    plain.load_state_dict(model.state_dict())
    plain.eval()
    with torch.no_grad():
        e3 = plain.native_engine().loss(ids, labels, None).item()
    assert abs(e1 - e3) < 1e-6


def test_engine_dropout_gradients_match_torch_with_same_masks() -> None:
    """One layer, the torch reference re-implements the block with the dumped masks."""
    import torch.nn.functional as F

    from llmtrain.models.gpt import GPT
    from llmtrain.ops import kernels as K

    torch.manual_seed(3)
    p = 0.2
    V, T, d, H, Fd, B = 256, 128, 128, 2, 512, 2
    model = GPT(vocab_size=V, block_size=T, d_model=d, n_layers=1, n_heads=H, d_ff=Fd, dropout=p,
                tie_embeddings=True).cuda()
    model.train()
    ids = torch.randint(0, V, (B, T), device="cuda")
    labels = torch.randint(0, V, (B, T), device="cuda")
    eng = model.native_engine()
    loss = eng.loss(ids, labels, None)
    loss.backward()
    seed, off = eng._dropout_seed, eng._dropout_calls
    scale = K.dropout_scale(p)
    M = B * T
    emb_mask = K.dropout_mask(M, d, p, seed=seed, offset=off, site=0xFFFF0000).float().view(B, T, d)
    att_mask = K.dropout_mask(B * H * T, T, p, seed=seed, offset=off, site=0).float()
    res_mask = K.dropout_mask(M, d, p, seed=seed, offset=off, site=1).float().view(B, T, d)
    mlp_mask = K.dropout_mask(M, d, p, seed=seed, offset=off, site=2).float().view(B, T, d)

    params = {k: v.detach().float().clone().requires_grad_(True) for k, v in model.named_parameters()}
    g = lambda name: params[name]  # noqa: E731
    x = (g("token_embedding.weight")[ids] + g("position_embedding.weight")[:T]) * emb_mask * scale
    blk = "blocks.0."
    h1 = F.layer_norm(x, (d,), g(blk + "ln_1.weight"), g(blk + "ln_1.bias"))
    qkv = h1 @ g(blk + "attn.qkv_proj.weight").T + g(blk + "attn.qkv_proj.bias")
    att = _ref_attention_dropout(qkv, H, att_mask, scale)
    x = x + (att @ g(blk + "attn.out_proj.weight").T + g(blk + "attn.out_proj.bias")) * res_mask * scale
    h2 = F.layer_norm(x, (d,), g(blk + "ln_2.weight"), g(blk + "ln_2.bias"))
    hid = F.gelu(h2 @ g(blk + "mlp_fc.weight").T + g(blk + "mlp_fc.bias"))
    x = x + (hid @ g(blk + "mlp_proj.weight").T + g(blk + "mlp_proj.bias")) * mlp_mask * scale
    xf = F.layer_norm(x, (d,), g("ln_f.weight"), g("ln_f.bias"))
    logits = xf @ g("token_embedding.weight").T
    ref_loss = F.cross_entropy(logits.view(-1, V), labels.view(-1))
    ref_loss.backward()
    assert abs(loss.item() - ref_loss.item()) < 3e-2
    worst = 0.0
    for name, prm in model.named_parameters():
        if prm.grad is None or name not in params:
            continue
        worst = max(worst, _rel(prm.grad, params[name].grad))
        assert _rel(prm.grad, params[name].grad) < 8e-2, name
    assert worst > 0.0
