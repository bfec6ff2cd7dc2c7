#!/usr/bin/env python
"""Does NCCL bring up NVLS (NVLink SHARP) on this box?  Run under torchrun with NCCL_DEBUG=INFO."""
import os

import torch
import torch.distributed as dist

torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
x = torch.ones(64 << 20, device="cuda")
for _ in range(3):
    dist.all_reduce(x)
torch.cuda.synchronize()
dist.destroy_process_group()
