#!/usr/bin/env python
"""Sample text (or raw token ids) from a checkpoint — the script form of the reference's
`notebooks/trained_vs_random_completion.ipynb`.

    python scripts/generate.py runs/<id>/checkpoints/step_000200.pt --prompt "The meaning of" \
        --max-new-tokens 40 --temperature 0.8 --top-k 40 [--device cuda] [--compare-random]
"""

from __future__ import annotations

import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "src"))


def main() -> int:
    import torch

    from llmtrain.generation import generate, load_model_from_checkpoint
    from llmtrain.models.gpt import GPTAdapter
    from llmtrain.models.tokenizer import UnavailableTokenizer

    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("checkpoint")
    ap.add_argument("--prompt", default=None, help="text prompt (needs the GPT-2 BPE tables)")
    ap.add_argument("--prompt-ids", default=None, help="comma separated token ids instead of text")
    ap.add_argument("--max-new-tokens", type=int, default=32)
    ap.add_argument("--temperature", type=float, default=0.8)
    ap.add_argument("--top-k", type=int, default=40)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--compare-random", action="store_true",
                    help="also sample from a freshly initialised model of the same shape")
    args = ap.parse_args()

    model, cfg = load_model_from_checkpoint(args.checkpoint, device=args.device)
    tok = GPTAdapter().build_tokenizer(cfg)
    have_tok = not isinstance(tok, UnavailableTokenizer)
    if args.prompt_ids:
        ids = [int(t) for t in args.prompt_ids.split(",")]
    elif args.prompt is not None and have_tok:
        ids = tok.encode(args.prompt)
    else:
        ids = [0]
    prompt = torch.tensor([ids], dtype=torch.long, device=args.device)
    print(f"model: {sum(p.numel() for p in model.parameters()):,} parameters on {args.device}")

    def sample(m, label: str) -> None:
        gen = torch.Generator(device=args.device).manual_seed(args.seed)
        out = generate(m, prompt, max_new_tokens=args.max_new_tokens,
                       temperature=args.temperature, top_k=args.top_k, generator=gen)[0].tolist()
        text = tok.decode(out) if have_tok and max(out) < tok.n_vocab else None
        print(f"--- {label}\n{text if text is not None else out}")

    sample(model, "trained")
    if args.compare_random:
        sample(GPTAdapter().build_model(cfg).to(args.device).eval(), "random init")
    return 0


if __name__ == "__main__":
    sys.exit(main())
