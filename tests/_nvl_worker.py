"""Worker for the multi-GPU NVLink tests (run under torch.distributed.run, one rank per GPU)."""

from __future__ import annotations

import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "src"))


def _check_allreduce(comm, rank: int, world: int) -> None:
    n = comm._n
    torch.manual_seed(100 + rank)
    src = torch.randn(n, device="cuda") * (rank + 1)
    out = torch.empty(n, device="cuda")
    for scale in (1.0, 1.0 / world):  # twice in a row: epochs / buffer reuse
        comm.allreduce(src, out, scale)
        torch.cuda.synchronize()
        rounded = src.to(torch.bfloat16).float()
        parts = [torch.empty_like(rounded) for _ in range(world)]
        dist.all_gather(parts, rounded)
        want = torch.zeros_like(rounded)
        for part in parts:  # rank order, fp32 — exactly what the kernel does
            want += part
        want *= scale
        assert torch.equal(out, want), f"rank {rank}: allreduce mismatch {(out - want).abs().max()}"


def _check_small_and_broadcast(comm, rank: int, world: int) -> None:
    for n in (8, 1000, 65536):  # twice each: the two send-buffer halves
        for rep in range(2):
            torch.manual_seed(50 + rank + rep)
            src = torch.randn(n, device="cuda")
            out = torch.empty(n, device="cuda")
            comm.allreduce_small(src, out, 0.5)
            torch.cuda.synchronize()
            rounded = src.to(torch.bfloat16).float()
            parts = [torch.empty_like(rounded) for _ in range(world)]
            dist.all_gather(parts, rounded)
            want = torch.zeros_like(rounded)
            for part in parts:
                want += part
            assert torch.equal(out, want * 0.5), f"one-shot mismatch n={n}"
    buf = torch.full((12345,), float(rank + 1), device="cuda")
    comm.broadcast(buf, root=world - 1)
    torch.cuda.synchronize()
    assert torch.equal(buf, torch.full_like(buf, float(world)))


def _stress_flags(comm, rank: int, world: int) -> None:
    """Back-to-back collectives with ranks arriving in a different order every time (random device
    sleeps): epoch flags must never be confused with an earlier round's (ABA), buffers never
    overwritten while a slow peer still reads them."""
    import random

    rng = random.Random(1234)  # same sequence on every rank
    n_small = 4096
    big_src = torch.empty(comm._n, device="cuda")
    big_out = torch.empty(comm._n, device="cuda")
    for it in range(40):
        late = rng.randrange(world)
        cycles = rng.choice((0, 200_000, 2_000_000))
        kind = rng.choice(("small", "big", "bcast"))
        if rank == late and cycles:
            torch.cuda._sleep(cycles)
        if kind == "small":
            src = torch.full((n_small,), float(rank + 1 + it), device="cuda")
            out = torch.empty_like(src)
            comm.allreduce_small(src, out, 1.0)
            want = float(sum(r + 1 + it for r in range(world)))
            torch.cuda.synchronize()
            assert torch.equal(out, torch.full_like(out, want)), f"stress small it={it}"
        elif kind == "big":
            big_src.fill_(float(rank + it % 7))
            comm.allreduce(big_src, big_out, 1.0)
            want = float(sum(r + it % 7 for r in range(world)))
            torch.cuda.synchronize()
            assert torch.equal(big_out, torch.full_like(big_out, want)), f"stress big it={it}"
        else:
            root = rng.randrange(world)
            buf = torch.full((777,), float(rank * 10 + it), device="cuda")
            comm.broadcast(buf, root=root)
            torch.cuda.synchronize()
            assert torch.equal(buf, torch.full_like(buf, float(root * 10 + it))), f"bcast it={it}"


def _check_step(comm, rank: int, world: int) -> None:
    from llmtrain.ops import kernels as K

    n = comm._n
    torch.manual_seed(7)
    p0 = torch.randn(n, device="cuda")
    p, m, v = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    p16 = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    norm = torch.zeros(1, device="cuda")
    ref_p, ref_m, ref_v = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    ref_p16 = torch.empty_like(p16)
    sumsq, ref_norm = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    for step in (1, 2, 3):
        torch.manual_seed(1000 * step + rank)
        g = torch.randn(n, device="cuda") * 0.01 * (rank + 1)
        hyper = dict(lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.1, step=step,
                     max_norm=1.0)
        # reference: gather everyone's bf16-rounded gradient, average, single-GPU fused AdamW
        rounded = g.to(torch.bfloat16).float()
        parts = [torch.empty_like(rounded) for _ in range(world)]
        dist.all_gather(parts, rounded)
        total = torch.zeros_like(rounded)
        for part in parts:
            total += part
        K.grad_sumsq(total, sumsq)
        K.adamw(ref_p, ref_m, ref_v, total, ref_p16, sumsq=sumsq, norm_out=ref_norm,
                grad_scale=1.0 / world, **hyper)
        comm.step(g, p, m, v, p16, norm, hyper)
        torch.cuda.synchronize()
        assert g.abs().max().item() == 0.0, "gradient arena was not zeroed"
        assert abs(norm.item() / ref_norm.item() - 1) < 1e-4, (norm.item(), ref_norm.item())
        assert (p - ref_p).abs().max().item() < 1e-5
        assert torch.equal(p16, p.to(torch.bfloat16))
    # replicas must be bit-identical
    mine = p.clone()
    others = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(others, mine)
    assert all(torch.equal(others[0], o) for o in others), "replicas diverged"


def _check_trainer(rank: int, world: int, overlap: bool = True, nvls: bool = False) -> None:
    from llmtrain.config.schemas import RunConfig
    from llmtrain.distributed import DDPState
    from llmtrain.training import Trainer

    cfg = RunConfig.model_validate({
        "run": {"name": "nvl-trainer", "device": "cuda"},
        "model": {"name": "gpt", "block_size": 128, "d_model": 128, "n_layers": 2, "n_heads": 2,
                  "d_ff": 512, "dropout": 0.0, "vocab_size": 1000},
        "data": {"name": "synthetic_text", "extra": {"pattern": "arith"}},
        "trainer": {"max_steps": 20, "micro_batch_size": 4, "grad_accum_steps": 2,
                    "warmup_steps": 0, "lr": 3e-3, "weight_decay": 0.0, "log_every_steps": 5,
                    "extra": {"overlap_comm": overlap}},
        "ddp": {"enabled": True, "backend": "nvl"},
        "mlflow": {"enabled": False}, "logging": {"log_to_file": False}, "output": {},
    })
    os.environ["LLMTRAIN_NVLS"] = "1" if nvls else "0"
    state = DDPState(rank=rank, world_size=world, local_rank=rank, is_main=rank == 0)
    torch.manual_seed(1234)  # same initial weights for every variant: their losses are comparable
    trainer = Trainer(cfg, ddp_state=state)
    comm = trainer._engine._comm
    assert comm is not None and comm.name.startswith("nvl"), getattr(comm, "name", None)
    assert comm.overlap is overlap
    if nvls and rank == 0:
        print("TRAINER_COMM", comm.name.split("(")[0], flush=True)
    result = trainer.fit()
    assert result.final_loss < result.first_step_loss - 0.3, (result.first_step_loss, result.final_loss)
    flat = trainer._engine._gpt.p32.clone()
    others = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(others, flat)
    assert all(torch.equal(others[0], o) for o in others), "replicas diverged after training"
    return result.final_loss


def _check_ddp_equivalence(rank: int, world: int, nvls: bool, backend: str = "nvl") -> None:
    """W ranks on their shards == one rank on the concatenated batch, from the same weights
    (the reference has no such test, `tests/test_distributed.py` only fakes the
    collectives).  Three optimiser steps; compared per step: the global mean loss and the
    gradient norm, and at the end the parameters (relative to the distance they travelled)."""
    from llmtrain.config.schemas import RunConfig
    from llmtrain.distributed import DDPState
    from llmtrain.training import Trainer

    mb, seq, steps = 4, 128, 3

    def cfg_for(batch: int, ddp: bool) -> RunConfig:
        return RunConfig.model_validate({
            "run": {"name": "ddp-eq", "device": "cuda"},
            "model": {"name": "gpt", "block_size": seq, "d_model": 128, "n_layers": 2,
                      "n_heads": 2, "d_ff": 512, "dropout": 0.0, "vocab_size": 1000},
            "data": {"name": "synthetic_text", "extra": {"pattern": "arith"}},
            "trainer": {"max_steps": steps, "micro_batch_size": batch, "grad_accum_steps": 1,
                        "warmup_steps": 0, "lr": 1e-3, "weight_decay": 0.0,
                        "extra": {"overlap_comm": True, "prefetch": False}},
            "ddp": {"enabled": ddp, "backend": backend, "timeout_sec": 60},
            "mlflow": {"enabled": False}, "logging": {"log_to_file": False}, "output": {},
        })

    def say(what: str) -> None:
        if os.environ.get("NVL_WORKER_VERBOSE"):
            print(f"[ddp-eq rank {rank}] {what}", flush=True)

    os.environ["LLMTRAIN_NVLS"] = "1" if nvls else "0"
    state = DDPState(rank=rank, world_size=world, local_rank=rank, is_main=rank == 0)
    torch.manual_seed(4321)
    many = Trainer(cfg_for(mb, True), ddp_state=state)._engine
    say("W-rank trainer built")
    torch.manual_seed(4321)
    one = Trainer(cfg_for(mb * world, False))._engine
    say("1-rank trainer built")
    one._gpt.p32.copy_(many._gpt.p32)
    one._gpt.refresh_shadow()
    start = many._gpt.p32.clone()
    gen = torch.Generator().manual_seed(99)  # identical on every rank
    for step in range(steps):
        tokens = torch.randint(0, 1000, (mb * world, seq + 1), generator=gen)
        whole = {"input_ids": tokens[:, :-1].contiguous(), "labels": tokens[:, 1:].contiguous()}
        shard = {k: v[rank * mb:(rank + 1) * mb].contiguous() for k, v in whole.items()}
        loss_many, _ = many.train_step(lambda shard=shard: shard, 1)
        say(f"step {step}: W-rank step enqueued")
        loss_one, _ = one.train_step(lambda whole=whole: whole, 1)
        torch.cuda.synchronize()
        say(f"step {step}: both steps done")
        mean = torch.tensor(float(loss_many), device="cuda")
        dist.all_reduce(mean)
        mean /= world
        assert abs(mean.item() - loss_one.item()) < 2e-3 * abs(loss_one.item()), (
            step, mean.item(), loss_one.item())
        n_many, n_one = many._norm.item(), one._norm.item()
        assert abs(n_many / n_one - 1) < 1e-2, (step, n_many, n_one)
    travelled = (one._gpt.p32 - start).pow(2).mean().sqrt().item()
    apart = (many._gpt.p32 - one._gpt.p32).pow(2).mean().sqrt().item()
    assert apart < 0.1 * travelled, (apart, travelled)
    if rank == 0:
        path = ("nccl" if backend == "nccl"
                else "nvls" if many._comm._team is not None else "p2p")
        print(f"DDP_EQUIVALENCE_OK world={world} path={path} rms_apart/rms_travelled="
              f"{apart / travelled:.4f}", flush=True)
    many._comm.close()


def _check_nvls(rank: int, world: int) -> None:
    """NVLink-SHARP all-reduce (multimem) against an fp32 sum of the gathered inputs."""
    from llmtrain.parallel.nvl import NvlComm, NvlsTeam, shutdown

    comm = NvlComm(1024 * 1024, rank=rank, world=world, timeout_s=60)
    n = 4 * 1024 * 1024
    team = NvlsTeam.try_create(comm, n)
    if team is None:
        if rank == 0:
            print("NVLS_UNAVAILABLE", flush=True)
        shutdown()
        return
    for count, scale in ((n, 1.0), (n // 2, 1.0 / world), (16 * world, 3.0)):
        torch.manual_seed(17 + rank)
        src = torch.randn(count, device="cuda") * (rank + 1)
        out = torch.empty(count, device="cuda")
        team.allreduce(src, out, scale=scale)
        torch.cuda.synchronize()
        parts = [torch.empty_like(src) for _ in range(world)]
        dist.all_gather(parts, src)
        want = torch.stack(parts).double().sum(0) * scale
        err = (out.double() - want).abs().max().item() / (want.abs().max().item() + 1e-12)
        assert err < 1e-5, f"rank {rank}: nvls allreduce rel err {err:.3e} at count {count}"
        others = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(others, out)
        assert all(torch.equal(others[0], o) for o in others), "ranks disagree on the result"
    # in place on the registered buffer, timed
    team.allreduce(None, None, count=n)
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(10):
        team.allreduce(None, None, count=n)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / 10
    if rank == 0:
        print(f"NVLS_OK {n * 4 / 1e6:.0f} MB in {ms * 1e3:.0f} us", flush=True)
    team.close()
    shutdown()


def main() -> int:
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("cpu:gloo,cuda:nccl")
    if len(sys.argv) > 1 and sys.argv[1] == "ddp-eq":
        try:
            from llmtrain.parallel.nvl import shutdown

            _check_ddp_equivalence(rank, world, nvls=False)
            _check_ddp_equivalence(rank, world, nvls=True)
            _check_ddp_equivalence(rank, world, nvls=False, backend="nccl")
            shutdown()
            dist.barrier()
            if rank == 0:
                print("NVL_WORKER_OK", flush=True)
        finally:
            dist.destroy_process_group()
        return 0
    if len(sys.argv) > 1 and sys.argv[1] == "nvls":
        import logging

        logging.basicConfig(level=logging.INFO if rank == 0 else logging.WARNING)
        try:
            _check_nvls(rank, world)
            dist.barrier()
            if rank == 0:
                print("NVL_WORKER_OK", flush=True)
        finally:
            dist.destroy_process_group()
        return 0
    try:
        from llmtrain.parallel.nvl import NvlComm, shutdown

        def note(what: str) -> None:
            if rank == 0:
                print(f"[nvl-worker] {what} ok", flush=True)

        comm = NvlComm(8 * 1024 * 1024 + 64, rank=rank, world=world, timeout_s=60)
        _check_allreduce(comm, rank, world)
        note("allreduce")
        _check_small_and_broadcast(comm, rank, world)
        note("one-shot + broadcast")
        _stress_flags(comm, rank, world)
        note("flag stress")
        _check_step(comm, rank, world)
        note("fused step")
        loss_overlapped = _check_trainer(rank, world, overlap=True)
        note("trainer overlapped")
        loss_serial = _check_trainer(rank, world, overlap=False)
        note("trainer serial")
        # in-switch bucket reduction (falls back to the peer-to-peer path where multicast is absent)
        loss_nvls = _check_trainer(rank, world, overlap=True, nvls=True)
        note("trainer nvls")
        # same seed, same data, same schedule: the three reduction paths train the same model
        assert abs(loss_nvls - loss_serial) < 0.05, (loss_nvls, loss_serial)
        assert abs(loss_overlapped - loss_serial) < 0.05, (loss_overlapped, loss_serial)
        _check_ddp_equivalence(rank, world, nvls=False)
        _check_ddp_equivalence(rank, world, nvls=True)
        _check_ddp_equivalence(rank, world, nvls=False, backend="nccl")
        shutdown()
        dist.barrier()
        if rank == 0:
            print("NVL_WORKER_OK")
    finally:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
