#!/usr/bin/env python
"""Gradient all-reduce sweep 1 KB - 1 GB (fp32 payload) — BASELINE config #5.

Run under torchrun with 2/4/8 ranks.  For each size: this repo's NVLink peer-memory all-reduce
(bf16 on the wire, fp32 accumulate, rank-ordered) vs NCCL ``all_reduce`` on fp32 as torch DDP
issues it.  Device-timed (CUDA events), max over ranks.  Reported bandwidths:

* ``nvls_*``: the NVLink-SHARP path (``multimem.ld_reduce`` / ``multimem.st``, fp32 on the wire,
  in place on a registered buffer), when multicast objects are available.
* ``algbw``  = payload bytes / time                      (payload = n * 4 bytes of fp32 gradient)
* ``busbw``  = algbw * 2 (W-1) / W                      (the NCCL-tests convention), vs 900 GB/s
  per direction per GPU nominal / 770 GB/s measured peer copy.
"""

from __future__ import annotations

import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def main() -> int:
    import torch
    import torch.distributed as dist

    from llmtrain.parallel.nvl import NvlComm, NvlsTeam

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("cpu:gloo,cuda:nccl")
    sizes = [1 << p for p in range(10, 31, 2)]  # bytes of fp32 payload: 1 KB .. 1 GB
    rows = []
    # NVLink-SHARP team over one registered buffer big enough for the largest payload (in-switch
    # reduction, fp32 on the wire); None where multicast objects cannot be created.
    anchor = NvlComm(1024, rank=rank, world=world, timeout_s=60)
    team = NvlsTeam.try_create(anchor, max(sizes) // 4)

    def timed(fn, iters: int) -> float:
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / iters], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for nbytes in sizes:
        n = max(8, nbytes // 4)
        n = (n + 7) // 8 * 8
        src = torch.randn(n, device="cuda")
        dst = torch.empty_like(src)
        comm = NvlComm(n, rank=rank, world=world, timeout_s=60)
        iters = 20 if nbytes <= (1 << 24) else 5
        ours = timed(lambda: comm.allreduce(src, dst, 1.0 / world), iters)
        buf = src.clone()
        nccl = timed(lambda: dist.all_reduce(buf, op=dist.ReduceOp.SUM), iters)
        # one-shot kernel (every rank reads all peers in ONE launch): the small-message algorithm,
        # available up to 262144 elements; the communicator's arena must hold 2 * n elements
        oneshot = None
        if n <= 262144:
            small = NvlComm(max(2 * n, 1024), rank=rank, world=world, timeout_s=60)
            oneshot = timed(lambda: small.allreduce_small(src, dst, 1.0 / world), iters)
            small.close()
        comm.close()
        nvls = None
        count = n // (4 * world) * (4 * world)
        if team is not None and count > 0:
            nvls = timed(lambda: team.allreduce(None, None, count=count, scale=1.0 / world), iters)
        factor = 2.0 * (world - 1) / world
        row = {
            "payload_bytes": n * 4, "world": world,
            "nvl_us": round(ours * 1e3, 2), "nccl_us": round(nccl * 1e3, 2),
            "nvl_busbw_GBs": round(n * 4 / ours / 1e6 * factor, 1),
            "nccl_busbw_GBs": round(n * 4 / nccl / 1e6 * factor, 1),
            "nvl_wire_bytes_in_per_gpu": int(n * 2 * (world - 1) / world + n * 4 * (world - 1) / world),
        }
        if oneshot is not None:
            row["oneshot_us"] = round(oneshot * 1e3, 2)
            row["oneshot_busbw_GBs"] = round(n * 4 / oneshot / 1e6 * factor, 1)
        if nvls is not None:
            row["nvls_us"] = round(nvls * 1e3, 2)
            row["nvls_busbw_GBs"] = round(count * 4 / nvls / 1e6 * factor, 1)
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if team is not None:
        team.close()
    anchor.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
