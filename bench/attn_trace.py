#!/usr/bin/env python
"""Per-tile clock64 timeline of one attention CTA (forward and backward); run with
LLMTRAIN_ATTN_TRACE=1.  Debug aid for the pipeline analysis in profiles/README.md."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "src"))
from llmtrain.ops import kernels as K  # noqa: E402

b, t, h = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (2, 1024, 2)))
qkv = torch.randn(b, t, 3 * h * 64, device="cuda").to(torch.bfloat16)
dout = torch.randn(b, t, h * 64, device="cuda").to(torch.bfloat16)
for _ in range(2):
    out, lse = K.attention_fwd(qkv, h)
    dqkv = K.attention_bwd(qkv, out, dout, lse, h)
torch.cuda.synchronize()
