"""CPU-side checks of the Python plumbing around the native runtime (no GPU, no library load)."""

from __future__ import annotations

import pytest


def test_supports_covers_reference_head_widths() -> None:
    """Every preset of the reference uses 16- or 32-wide heads; GPT-2 uses 64."""
    from llmtrain.engine.b200 import supports
    from llmtrain.models.gpt import GPT

    def model(d_model: int, n_heads: int) -> GPT:
        return GPT(vocab_size=32, block_size=8, d_model=d_model, n_layers=1, n_heads=n_heads,
                   d_ff=64, dropout=0.1)

    assert supports(model(64, 4))     # head_dim 16
    assert supports(model(256, 8))    # head_dim 32
    assert supports(model(128, 2))    # head_dim 64
    assert not supports(model(256, 2))  # head_dim 128: torch-op path
    assert not supports(model(72, 6))   # head_dim 12 is not a multiple of 8
    assert not supports(object())       # type: ignore[arg-type]


def test_dropout_scale_is_the_exact_inverse_of_the_quantised_keep_probability() -> None:
    from llmtrain.ops.kernels import dropout_scale

    assert dropout_scale(0.0) == 1.0
    for p in (0.1, 0.2, 0.5, 0.9):
        threshold = min(max(int(p * 65536.0 + 0.5), 1), 65535)
        keep = 1.0 - threshold / 65536.0
        assert abs(dropout_scale(p) * keep - 1.0) < 1e-12
        assert abs(keep - (1.0 - p)) < 1.0 / 65536.0


@pytest.mark.parametrize(
    "env,world,overlap,expect",
    [(None, 2, True, False), (None, 4, True, True), (None, 8, True, True), (None, 8, False, False),
     ("0", 8, True, False), ("1", 2, True, True), ("1", 2, False, False)],
)
def test_nvls_selection_policy(monkeypatch, env, world, overlap, expect) -> None:
    """In-switch reduction from 4 GPUs up, only on the overlapped path, env var overrides."""
    from llmtrain.parallel import nvl

    if env is None:
        monkeypatch.delenv("LLMTRAIN_NVLS", raising=False)
    else:
        monkeypatch.setenv("LLMTRAIN_NVLS", env)
    assert nvl._want_nvls(world, overlap) is expect


def test_native_library_path_can_be_overridden(monkeypatch, tmp_path) -> None:
    import importlib

    fake = tmp_path / "libother.so"
    monkeypatch.setenv("LLMTRAIN_NATIVE_LIB", str(fake))
    import llmtrain.ops.native as native

    reloaded = importlib.reload(native)
    try:
        assert reloaded.library_path() == fake
    finally:
        monkeypatch.delenv("LLMTRAIN_NATIVE_LIB")
        importlib.reload(native)
