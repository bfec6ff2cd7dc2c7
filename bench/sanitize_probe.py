#!/usr/bin/env python
"""Launch every native kernel once on tiny shapes: the target of `scripts/sanitize.sh`
(compute-sanitizer memcheck / racecheck / synccheck).  Exits non-zero on a numerical mismatch so a
sanitizer run also re-checks results."""

from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def main() -> int:
    import torch

    from llmtrain.ops import gemm as G
    from llmtrain.ops import kernels as K

    torch.manual_seed(0)
    dev = "cuda"

    def bf(*shape, s=1.0):
        return (torch.randn(*shape, device=dev) * s).to(torch.bfloat16)

    def close(a, b, tol):
        err = ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()
        assert err < tol, err

    # GEMM family (CTA pairs for M > 128, single CTA otherwise)
    for m in (128, 320):
        n, k = 192, 128
        x, w, bias = bf(m, k), bf(n, k, s=0.05), torch.randn(n, device=dev)
        out, out2, r = (torch.empty(m, n, device=dev, dtype=torch.bfloat16) for _ in range(3))
        r.copy_(bf(m, n))
        G.gemm(G.NT, G.EPI_BIAS, x, w, out, bias=bias)
        close(out, x.float() @ w.float().T + bias, 2e-2)
        G.gemm(G.NT, G.EPI_BIAS_GELU, x, w, out, c2=out2, bias=bias)
        G.gemm(G.NT, G.EPI_BIAS_RESID, x, w, out, bias=bias, r=r)
        close(out, x.float() @ w.float().T + bias + r.float(), 2e-2)
        dy, wt = bf(m, n), bf(n, k, s=0.05)
        dx = torch.empty(m, k, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NN, G.EPI_BIAS, dy, wt, dx)
        close(dx, dy.float() @ wt.float(), 2e-2)
        rk, cs = bf(m, k), torch.zeros(k, device=dev)
        G.gemm_colsum(G.NN, G.EPI_DGELU, dy, wt, dx, cs, r=rk)
        close(cs, dx.float().sum(0), 1e-2)
        dw = torch.zeros(n, k, device=dev)
        G.gemm(G.TN, G.EPI_ACCUM_F32, dy, x, dw)
        close(dw, dy.float().T @ x.float(), 2e-2)
        K.linear_dropout_resid(x, w, bias, r, 0.2, seed=1, offset=2, site=3, out=out)

    # attention (plain, dropout, key padding)
    qkv = bf(2, 200, 3 * 128)
    mask = torch.ones(2, 200, dtype=torch.uint8, device=dev)
    mask[1, 150:] = 0
    for kw in ({}, {"dropout": (0.1, 5, 6, 7)}, {"key_mask": mask}):
        o, lse = K.attention_fwd(qkv, 2, **kw)
        K.attention_bwd(qkv, o, bf(2, 200, 128), lse, 2, **kw)

    # LayerNorm, embedding, cross-entropy, bias grad, optimiser, dropout, head padding
    x = bf(70, 768)
    gamma, beta = torch.rand(768, device=dev) + 0.5, torch.randn(768, device=dev)
    mean, rstd = torch.empty(70, device=dev), torch.empty(70, device=dev)
    y = K.layernorm_fwd(x, gamma, beta, mean=mean, rstd=rstd)
    y = y[0] if isinstance(y, tuple) else y
    close(y, torch.nn.functional.layer_norm(x.float(), (768,), gamma, beta), 2e-2)
    dg, db, dc = (torch.zeros(768, device=dev) for _ in range(3))
    K.layernorm_bwd(bf(70, 768), x, gamma, mean, rstd, bf(70, 768), dg, db, dcolsum=dc)
    ids = torch.randint(0, 500, (2, 35), device=dev)
    wte, wpe = bf(500, 64), bf(40, 64)
    K.embed_fwd(ids, wte, wpe)
    K.embed_bwd(ids, bf(2, 35, 64).view(70, 64), torch.zeros(500, 64, device=dev),
                torch.zeros(40, 64, device=dev))
    logits = bf(33, 1024, s=2.0)
    K.cross_entropy(logits, torch.randint(0, 1000, (33,), device=dev), vocab=1000,
                    loss_acc=torch.zeros(1, device=dev), grad_scale=1.0 / 33)
    K.bias_grad(bf(300, 96), torch.zeros(96, device=dev))
    n = 10_000
    p, g = torch.randn(n, device=dev), torch.randn(n, device=dev)
    m_, v_ = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    p16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
    sumsq, norm = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
    K.grad_sumsq(g, sumsq)
    K.adamw(p, m_, v_, g, p16, sumsq=sumsq, norm_out=norm, lr=1e-3, beta1=0.9, beta2=0.999,
            eps=1e-8, weight_decay=0.1, step=1, max_norm=1.0)
    K.dropout(bf(37, 50), 0.3, seed=1, offset=2, site=3)
    K.dropout_mask(37, 50, 0.3, seed=1, offset=2, site=3)
    torch.cuda.synchronize()
    print("sanitize probe ok")
    return 0


if __name__ == "__main__":
    sys.exit(main())
