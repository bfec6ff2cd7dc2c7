"""tcgen05 GEMM vs an fp32 torch reference of the same op (run on the B200 box)."""

from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, scale=1.0, dtype=torch.bfloat16):
    return (torch.randn(*shape, device="cuda", dtype=torch.float32) * scale).to(dtype)


def _close(got: torch.Tensor, want: torch.Tensor, tol: float) -> None:
    got = got.float()
    err = (got - want).abs().max().item()
    ref = want.abs().max().item() + 1e-6
    assert err / ref < tol, f"max abs err {err:.4e} vs ref scale {ref:.4e}"


SHAPES = [(256, 256, 128), (384, 512, 768), (1000, 320, 1600), (128, 2304, 64), (4096, 768, 3072)]


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_nt_bias(m: int, n: int, k: int) -> None:
    from llmtrain.ops import gemm as G

    x, w = _rand((m, k)), _rand((n, k), 0.05)
    bias = torch.randn(n, device="cuda")
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NT, G.EPI_BIAS, x, w, out, bias=bias)
    _close(out, x.float() @ w.float().t() + bias, 1e-2)
    out2 = torch.empty_like(out)
    G.gemm(G.NT, G.EPI_BIAS, x, w, out2)  # no bias
    _close(out2, x.float() @ w.float().t(), 1e-2)


@pytest.mark.parametrize("m,n,k", SHAPES[:3])
def test_nt_gelu_and_resid(m: int, n: int, k: int) -> None:
    from llmtrain.ops import gemm as G

    x, w = _rand((m, k)), _rand((n, k), 0.05)
    bias = torch.randn(n, device="cuda") * 0.1
    dact = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    act = torch.empty_like(dact)
    G.gemm(G.NT, G.EPI_BIAS_GELU, x, w, act, c2=dact, bias=bias)
    want_pre = (x.float() @ w.float().t() + bias).requires_grad_(True)
    want_act = torch.nn.functional.gelu(want_pre)  # exact (erf) GELU
    want_act.sum().backward()
    _close(act, want_act.detach(), 1e-2)
    _close(dact, want_pre.grad, 1e-2)  # second output: gelu'(pre), consumed by the MLP dgrad
    want_pre = want_pre.detach()

    resid = _rand((m, n))
    out = torch.empty_like(act)
    G.gemm(G.NT, G.EPI_BIAS_RESID, x, w, out, bias=bias, r=resid)
    _close(out, want_pre + resid.float(), 1e-2)


@pytest.mark.parametrize("m,n,k", SHAPES[:4])
def test_nn_dgrad(m: int, n: int, k: int) -> None:
    from llmtrain.ops import gemm as G

    dy, w = _rand((m, k)), _rand((k, n), 0.05)  # C[M,N] = dy[M,K] @ w[K,N]
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NN, G.EPI_BIAS, dy, w, out)
    want = dy.float() @ w.float()
    _close(out, want, 1e-2)

    factor = _rand((m, n))  # gelu'(pre) as saved by the forward GELU epilogue
    out2 = torch.empty_like(out)
    G.gemm(G.NN, G.EPI_DGELU, dy, w, out2, r=factor)
    _close(out2, want * factor.float(), 1e-2)


@pytest.mark.parametrize("m,n,k", [(256, 256, 512), (768, 2304, 4096), (320, 1600, 1000), (50304, 768, 2048)])
def test_tn_wgrad_accumulates(m: int, n: int, k: int) -> None:
    from llmtrain.ops import gemm as G

    dy, x = _rand((k, m), 0.1), _rand((k, n))  # C[M,N] += dy^T @ x
    base = torch.randn(m, n, device="cuda")
    out = base.clone()
    G.gemm(G.TN, G.EPI_ACCUM_F32, dy, x, out)
    _close(out, base + dy.float().t() @ x.float(), 5e-3)


def test_strided_views() -> None:
    """Operands may be column slices of a wider matrix (q/k/v inside the fused qkv buffer)."""
    from llmtrain.ops import gemm as G

    wide = _rand((512, 3 * 256))
    a = wide[:, 256:512]
    w = _rand((384, 256), 0.05)
    out_wide = torch.zeros(512, 1024, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NT, G.EPI_BIAS, a, w, out_wide[:, 128:512])
    _close(out_wide[:, 128:512], a.float() @ w.float().t(), 1e-2)
    assert out_wide[:, :128].abs().max().item() == 0 and out_wide[:, 512:].abs().max().item() == 0


@pytest.mark.parametrize("m,n,k", [(512, 3072, 768), (1000, 320, 1600), (130, 72, 64)])
def test_dgrad_with_fused_column_sums(m: int, n: int, k: int) -> None:
    """dX = (dY @ W) * R with colsum += dX.sum(0) out of the same epilogue (fc bias gradient)."""
    from llmtrain.ops import gemm as G

    dy, w = _rand((m, k)), _rand((k, n), 0.05)
    r = _rand((m, n))
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    colsum = torch.full((n,), 2.0, device="cuda")
    G.gemm_colsum(G.NN, G.EPI_DGELU, dy, w, out, colsum, r=r)
    want = (dy.float() @ w.float()) * r.float()
    _close(out, want, 2e-2)
    # the sums are taken over the bf16 values that were stored
    got, ref = colsum - 2.0, out.float().sum(0)
    assert ((got - ref).abs().max() / (ref.abs().max() + 1e-6)).item() < 1e-3


_ALT_PATH_CHECK = r"""
import torch
from llmtrain.ops import gemm as G
torch.manual_seed(0)
m, n, k = 640, 512, 384
x = (torch.randn(m, k, device="cuda")).to(torch.bfloat16)
w = (torch.randn(n, k, device="cuda") * 0.05).to(torch.bfloat16)
bias = torch.randn(n, device="cuda")
out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
G.gemm(G.NT, G.EPI_BIAS, x, w, out, bias=bias)
want = x.float() @ w.float().T + bias
err = ((out.float() - want).abs().max() / want.abs().max()).item()
dw = torch.zeros(n, k, device="cuda")
dy = (torch.randn(m, n, device="cuda") * 0.1).to(torch.bfloat16)
G.gemm(G.TN, G.EPI_ACCUM_F32, dy, x, dw)
want_dw = dy.float().T @ x.float()
err2 = ((dw - want_dw).abs().max() / want_dw.abs().max()).item()
assert err < 2e-2 and err2 < 2e-2, (err, err2)
print("ok")
"""


@pytest.mark.parametrize("env", [{"LLMTRAIN_GEMM_PAIR": "0"}, {"LLMTRAIN_PDL": "0"},
                                 {"LLMTRAIN_GEMM_GROUP_M": "2"}])
def test_alternative_launch_paths_in_a_fresh_process(env: dict) -> None:
    """The single-CTA kernel, plain stream order (no PDL) and narrow tile bands are selected by
    environment variables that the library reads once, so each is exercised in its own process."""
    import os
    import subprocess
    import sys
    from pathlib import Path

    src = Path(__file__).resolve().parents[1] / "src"
    proc = subprocess.run(
        [sys.executable, "-c", _ALT_PATH_CHECK],
        env={**os.environ, **env, "PYTHONPATH": str(src)},
        capture_output=True, text=True, timeout=300,
    )
    assert proc.returncode == 0, proc.stderr[-2000:]
    assert "ok" in proc.stdout
