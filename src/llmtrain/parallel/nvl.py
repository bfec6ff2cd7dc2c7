"""Gradient reduction + optimiser step across ranks.

* :class:`NvlComm` — this repo's NVLink / NVSwitch peer-memory path (``csrc/comm/nvl.cu``): CUDA-IPC
  mapped symmetric buffers, flag barriers, rank-ordered reduce-scatter, and an all-gather that is
  fused with average + global-norm clip + AdamW + bf16 shadow refresh.  It replaces the c10d
  Reducer + ProcessGroup all-reduce the reference rents through ``DistributedDataParallel``
  (`src/llmtrain/training/trainer.py:87-91,377-394`).  The torch process group is used only to
  exchange the IPC handles.
* :class:`NcclFallback` — the transport across nodes, or whenever peers cannot map each other (pods
  without a shared IPC namespace, no P2P): one NCCL all-reduce over the flat fp32 gradient arena
  followed by the same fused optimiser kernel.
"""

from __future__ import annotations

import ctypes
import logging
from typing import Any

import torch
import torch.distributed as dist

from llmtrain.ops import kernels as K
from llmtrain.ops import native
from llmtrain.ops.native import check, ptr, stream_handle

logger = logging.getLogger(__name__)

_IPC_HANDLE_BYTES = 64
_HANDLES_PER_RANK = 4  # send, red, sync, full
_live: list[Any] = []


def _collective_stage(what: str, fn: Any, world: int, on_fail: Any = None) -> None:
    """Run one set-up step on every rank; if ANY rank failed, every rank raises together.

    Set-up steps can fail on one rank only (an allocation, an IPC export, a handle import).  A
    rank that raised on its own would leave the others inside the next collective until
    ``ddp.timeout_sec`` expires; gathering the error strings after every step keeps the ranks in
    lock-step, so a failure becomes a clean fallback to NCCL instead of a hang.
    """
    err = ""
    try:
        fn()
    except Exception as exc:  # noqa: BLE001  (re-raised collectively below)
        err = f"{what}: {exc}"
    seen: list[Any] = [None] * world
    dist.all_gather_object(seen, err)
    bad = [e for e in seen if e]
    if bad:
        if on_fail is not None:
            on_fail()
        raise RuntimeError(bad[0])


_BUCKET_HOOK = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t)


class NcclFallback:
    """torch.distributed (NCCL) transport: the path across nodes and between pods that cannot map
    each other's memory.

    ``overlap=True`` (default): the last micro-batch's backward reports every gradient bucket as
    soon as its kernels are enqueued (``llmtrain_b200_gpt_backward_hooked``) and one asynchronous
    fp32 all-reduce per bucket starts on a side stream while backward continues — the job the
    c10d Reducer does for the reference (`src/llmtrain/training/trainer.py:88-91,377-387`), with
    this engine's per-layer buckets instead of 25 MiB ones.  The fused clip + AdamW kernel follows
    once the last bucket has landed.  ``overlap=False``: one all-reduce of the whole arena after
    backward.
    """

    def __init__(self, gpt: Any, *, overlap: bool = True) -> None:
        self._gpt = gpt
        self._world = dist.get_world_size()
        self.overlap = overlap
        self.name = ("nccl(per-layer fp32 buckets overlapped with backward -> local adamw)"
                     if overlap else "nccl(one fp32 all-reduce after backward -> local adamw)")
        self._side = torch.cuda.Stream() if overlap else None
        self._works: list[Any] = []
        self._reduced = False

    def backward_last(self) -> None:
        g = self._gpt
        if not self.overlap:
            g.run_backward()
            return
        arena = g.g32
        compute = torch.cuda.current_stream()
        side = self._side
        works = self._works
        works.clear()

        failure: list[BaseException] = []

        def on_bucket(_ctx: Any, lo: int, length: int) -> int:
            # called from C++ between kernel launches: an exception must not cross that frame
            try:
                if length == 0:
                    return 0
                ready = torch.cuda.Event()
                ready.record(compute)
                side.wait_event(ready)
                with torch.cuda.stream(side):
                    works.append(dist.all_reduce(arena[lo:lo + length], op=dist.ReduceOp.SUM,
                                                 async_op=True))
                return 0
            except BaseException as exc:  # noqa: BLE001  (re-raised below, on the Python side)
                failure.append(exc)
                return 1

        hook = _BUCKET_HOOK(on_bucket)
        lib = native.lib()
        lib.llmtrain_b200_gpt_backward_hooked.argtypes = [ctypes.c_void_p, _BUCKET_HOOK,
                                                          ctypes.c_void_p, ctypes.c_void_p]
        status = lib.llmtrain_b200_gpt_backward_hooked(g._handle, hook, None, stream_handle())
        if failure:
            raise failure[0]
        check(status, "gpt_backward_hooked")
        self._reduced = True

    def reduce_and_step(self, m: torch.Tensor, v: torch.Tensor, sumsq: torch.Tensor,
                        norm: torch.Tensor, hyper: dict[str, Any]) -> None:
        g = self._gpt
        if self._reduced:
            for work in self._works:
                work.wait()  # the compute stream waits for that bucket's reduction
            self._works.clear()
            self._reduced = False
        else:
            dist.all_reduce(g.g32, op=dist.ReduceOp.SUM)
        K.grad_sumsq(g.g32, sumsq)
        K.adamw(g.p32, m, v, g.g32, g.p16, sumsq=sumsq, norm_out=norm,
                grad_scale=1.0 / self._world, **hyper)

    def reduced_gradient(self) -> torch.Tensor:
        return self._gpt.g32

    def close(self) -> None:
        return None


def _peers_reachable(world: int) -> bool:
    """All ranks on this host, each with P2P access to every other visible device?"""
    count = torch.cuda.device_count()
    if count < world:
        return False
    me = torch.cuda.current_device()
    return all(d == me or torch.cuda.can_device_access_peer(me, d) for d in range(world))


class NvlComm:
    """Peer-memory communicator over an arena of ``n`` fp32 elements."""

    name = "nvl-p2p(pack-bf16 -> reduce-scatter -> gather+adamw)"

    def __init__(self, n: int, *, rank: int, world: int, timeout_s: float) -> None:
        self._lib = native.lib()
        self._n = n
        self.rank, self.world = rank, world
        self._handle = ctypes.c_void_p()
        handle = ctypes.c_void_p()

        def create() -> None:
            check(
                self._lib.llmtrain_b200_nvl_create(
                    ctypes.c_int(rank), ctypes.c_int(world), ctypes.c_size_t(n),
                    ctypes.c_double(float(timeout_s)), ctypes.byref(handle),
                ),
                "nvl_create",
            )
            self._handle = handle

        _collective_stage("nvl_create", create, world, self.close)
        mine = (ctypes.c_ubyte * (_HANDLES_PER_RANK * _IPC_HANDLE_BYTES))()
        _collective_stage(
            "nvl_export",
            lambda: check(self._lib.llmtrain_b200_nvl_export(handle, mine), "nvl_export"),
            world, self.close)
        gathered: list[bytes | None] = [None] * world
        dist.all_gather_object(gathered, bytes(mine))
        blob = b"".join(g for g in gathered if g is not None)

        def connect() -> None:
            if len(blob) != world * _HANDLES_PER_RANK * _IPC_HANDLE_BYTES:
                raise RuntimeError("nvl: IPC handle exchange returned an unexpected payload size")
            everyone = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
            check(self._lib.llmtrain_b200_nvl_connect(handle, everyone), "nvl_connect")

        # (the gather inside the stage doubles as the barrier "every rank has mapped every peer
        # before the first kernel runs")
        _collective_stage("nvl_connect", connect, world, self.close)
        _live.append(self)

    def barrier(self) -> None:
        check(self._lib.llmtrain_b200_nvl_barrier(self._handle, stream_handle()), "nvl_barrier")

    def allreduce(self, src: torch.Tensor, dst: torch.Tensor, scale: float = 1.0) -> None:
        """``dst = scale * sum_r bf16(src_r)`` (fp32 in/out, bf16 on the wire)."""
        assert src.numel() == self._n and dst.numel() == self._n
        check(
            self._lib.llmtrain_b200_nvl_allreduce(self._handle, ptr(src), ptr(dst),
                                                  ctypes.c_float(scale), stream_handle()),
            "nvl_allreduce",
        )

    def allreduce_small(self, src: torch.Tensor, dst: torch.Tensor, scale: float = 1.0) -> None:
        """One-launch all-reduce for messages up to 262144 elements (every rank reads all peers)."""
        check(
            self._lib.llmtrain_b200_nvl_allreduce_oneshot(
                self._handle, ptr(src), ptr(dst), ctypes.c_size_t(src.numel()),
                ctypes.c_float(scale), stream_handle()),
            "nvl_allreduce_oneshot",
        )

    def broadcast(self, buf: torch.Tensor, root: int = 0) -> None:
        """Copy-engine broadcast of an fp32 buffer (at most the arena length) from ``root``."""
        check(
            self._lib.llmtrain_b200_nvl_broadcast(self._handle, ptr(buf),
                                                  ctypes.c_size_t(buf.numel()),
                                                  ctypes.c_int(root), stream_handle()),
            "nvl_broadcast",
        )

    def step(self, g: torch.Tensor, p: torch.Tensor, m: torch.Tensor, v: torch.Tensor,
             p16: torch.Tensor, norm: torch.Tensor | None, hyper: dict[str, Any]) -> None:
        check(
            self._lib.llmtrain_b200_nvl_step(
                self._handle, ptr(g), ptr(p), ptr(m), ptr(v), ptr(p16), ptr(norm),
                ctypes.c_float(hyper["lr"]), ctypes.c_float(hyper["beta1"]),
                ctypes.c_float(hyper["beta2"]), ctypes.c_float(hyper["eps"]),
                ctypes.c_float(hyper["weight_decay"]), ctypes.c_int(hyper["step"]),
                ctypes.c_float(1.0 / self.world), ctypes.c_float(hyper["max_norm"]),
                stream_handle(),
            ),
            "nvl_step",
        )

    def backward_overlapped(self, gpt_handle: ctypes.c_void_p, g: torch.Tensor) -> None:
        """Backward of the bound GPT runtime; each gradient bucket is reduced across ranks on the
        communicator's own stream as soon as backward has finalised it."""
        check(
            self._lib.llmtrain_b200_nvl_backward_overlapped(self._handle, gpt_handle, ptr(g),
                                                            stream_handle()),
            "nvl_backward_overlapped",
        )

    def finish_step(self, g: torch.Tensor, p: torch.Tensor, m: torch.Tensor, v: torch.Tensor,
                    p16: torch.Tensor, norm: torch.Tensor | None, hyper: dict[str, Any]) -> None:
        check(
            self._lib.llmtrain_b200_nvl_finish_step(
                self._handle, ptr(g), ptr(p), ptr(m), ptr(v), ptr(p16), ptr(norm),
                ctypes.c_float(hyper["lr"]), ctypes.c_float(hyper["beta1"]),
                ctypes.c_float(hyper["beta2"]), ctypes.c_float(hyper["eps"]),
                ctypes.c_float(hyper["weight_decay"]), ctypes.c_int(hyper["step"]),
                ctypes.c_float(1.0 / self.world), ctypes.c_float(hyper["max_norm"]),
                stream_handle(),
            ),
            "nvl_finish_step",
        )

    def close(self) -> None:
        if self._handle:
            self._lib.llmtrain_b200_nvl_destroy.argtypes = [ctypes.c_void_p]
            self._lib.llmtrain_b200_nvl_destroy.restype = None
            self._lib.llmtrain_b200_nvl_destroy(self._handle)
            self._handle = ctypes.c_void_p()


class NvlsTeam:
    """NVLink-SHARP all-reduce over a registered fp32 buffer (``csrc/comm/nvls.cu``).

    The NVSwitch adds the ranks' values in flight: every rank pulls its 1/W slice with
    ``multimem.ld_reduce`` and pushes the result with ``multimem.st``.  Needs a device that reports
    multicast support; :func:`NvlsTeam.try_create` returns ``None`` (with the reason logged) when
    the driver or the fabric does not offer it, so callers keep using :class:`NvlComm`.
    """

    def __init__(self, comm: NvlComm, n: int, *, use_fabric: bool = True) -> None:
        self._lib = native.lib()
        self._lib.llmtrain_b200_nvls_buffer.restype = ctypes.c_void_p
        self._comm = comm
        self.n = n
        self._handle = ctypes.c_void_p()
        world = comm.world

        def stage(what: str, fn: Any) -> None:
            _collective_stage(what, fn, world, self.close)

        handle = ctypes.c_void_p()
        stage("nvls_create", lambda: check(
            self._lib.llmtrain_b200_nvls_create(
                comm._handle, ctypes.c_int(comm.rank), ctypes.c_int(world), ctypes.c_size_t(n),
                ctypes.c_int(1 if use_fabric else 0), ctypes.byref(handle)),
            "nvls_create"))
        self._handle = handle
        wire = (ctypes.c_ubyte * 80)()
        if comm.rank == 0:
            stage("nvls_export",
                  lambda: check(self._lib.llmtrain_b200_nvls_export(handle, wire), "nvls_export"))
        else:
            stage("nvls_export", lambda: None)
        payload: list[Any] = [bytes(wire) if comm.rank == 0 else None]
        dist.broadcast_object_list(payload, src=0)
        shared = (ctypes.c_ubyte * 80).from_buffer_copy(payload[0])
        # (the all-gather inside stage() doubles as the host barrier the driver API asks for:
        # every device has joined the multicast object before anyone binds memory to it)
        stage("nvls_connect",
              lambda: check(self._lib.llmtrain_b200_nvls_connect(handle, shared), "nvls_connect"))
        stage("nvls_bind", lambda: check(self._lib.llmtrain_b200_nvls_bind(handle), "nvls_bind"))

    @classmethod
    def try_create(cls, comm: NvlComm, n: int) -> NvlsTeam | None:
        lib = native.lib()
        if not lib.llmtrain_b200_nvls_supported():
            logger.info("nvls: multicast is not supported on this device/driver")
            return None
        # one node: a POSIX file descriptor (picked up with pidfd_getfd) is the natural handle;
        # fabric handles need the IMEX service and are only tried second
        for use_fabric in (False, True):
            try:
                return cls(comm, n, use_fabric=use_fabric)
            except RuntimeError as exc:  # raised on every rank together, see stage()
                logger.info("nvls: %s handles unavailable (%s)",
                            "fabric" if use_fabric else "posix-fd", exc)
        return None

    def allreduce(self, src: torch.Tensor | None, dst: torch.Tensor | None, *,
                  count: int | None = None, scale: float = 1.0) -> None:
        """``dst = scale * sum_r src_r`` (fp32).  ``src``/``dst`` ``None``: in place on the buffer."""
        n = count if count is not None else (src.numel() if src is not None else self.n)
        check(
            self._lib.llmtrain_b200_nvls_allreduce(
                self._handle, ptr(src), ptr(dst), ctypes.c_size_t(n), ctypes.c_float(scale),
                stream_handle(),
            ),
            "nvls_allreduce",
        )

    def buffer_tensor(self, n: int | None = None) -> torch.Tensor:
        """fp32 tensor aliasing this rank's view of the registered buffer (no copy)."""
        n = self.n if n is None else n
        assert 0 < n <= self.n
        self._lib.llmtrain_b200_nvls_buffer.restype = ctypes.c_void_p
        address = int(self._lib.llmtrain_b200_nvls_buffer(self._handle))

        class _Raw:  # torch aliases device memory described by the CUDA array interface
            __cuda_array_interface__ = {
                "shape": (n,), "typestr": "<f4", "data": (address, False), "version": 3,
                "strides": None,
            }

        raw = _Raw()
        tensor = torch.as_tensor(raw, device=torch.device("cuda", torch.cuda.current_device()))
        tensor._nvls_owner = (self, raw)  # type: ignore[attr-defined]  # keep the mapping alive
        return tensor

    def attach_to(self, comm: NvlComm) -> None:
        """Put ``comm``'s overlapped bucket path into in-switch mode on this buffer."""
        self._lib.llmtrain_b200_nvls_buffer.restype = ctypes.c_void_p
        self._lib.llmtrain_b200_nvls_mc_buffer.restype = ctypes.c_void_p
        uc = ctypes.c_void_p(self._lib.llmtrain_b200_nvls_buffer(self._handle))
        mc = ctypes.c_void_p(self._lib.llmtrain_b200_nvls_mc_buffer(self._handle))
        check(self._lib.llmtrain_b200_nvl_attach_nvls(comm._handle, uc, mc), "nvl_attach_nvls")

    def close(self) -> None:
        if self._handle:
            self._lib.llmtrain_b200_nvls_destroy(self._handle)
            self._handle = ctypes.c_void_p()


class _NvlTrainComm:
    """Adapter with the face the train engine expects.

    ``overlap=True`` (default): the last micro-batch's backward goes through
    :meth:`backward_last`, which reduces every per-layer bucket on a side stream while backward
    continues (owners push reduced slices to all peers), so ``reduce_and_step`` only has to run the
    local clip + AdamW pass.  ``overlap=False``: one pack / reduce-scatter / gather+AdamW sequence
    after backward.
    """

    def __init__(self, gpt: Any, comm: NvlComm, *, overlap: bool = True,
                 team: NvlsTeam | None = None) -> None:
        self._gpt, self._comm = gpt, comm
        # shutdown() must release the multicast team (object, fd, mappings) as well as the peer
        # mappings: this adapter takes the inner communicator's place in the live list
        if comm in _live:
            _live.remove(comm)
        _live.append(self)
        self.overlap = overlap
        self._reduced = False
        self._team = team
        self.name = ("nvl-p2p(per-layer buckets overlapped with backward: pack-bf16 -> "
                     "reduce+push -> local adamw)") if overlap else comm.name
        if team is not None:
            # NVLink-SHARP mode: the gradient arena IS the multicast-bound buffer, each bucket is
            # summed inside the switch, in place (multimem.ld_reduce / multimem.st).
            gpt.g32 = team.buffer_tensor(gpt.n_params)
            gpt.g32.zero_()
            gpt._bind()
            team.attach_to(comm)
            self.name = ("nvls(per-layer buckets overlapped with backward: in-switch "
                         "multimem.ld_reduce/st in place on the fp32 gradient arena -> local adamw)")

    def backward_last(self) -> None:
        g = self._gpt
        if self.overlap:
            self._comm.backward_overlapped(g._handle, g.g32)
            self._reduced = True
        else:
            g.run_backward()

    def reduce_and_step(self, m: torch.Tensor, v: torch.Tensor, sumsq: torch.Tensor,
                        norm: torch.Tensor, hyper: dict[str, Any]) -> None:
        del sumsq  # the global norm is assembled from the owners' slice partials on the device
        g = self._gpt
        if self._reduced:
            self._comm.finish_step(g.g32, g.p32, m, v, g.p16, norm, hyper)
            self._reduced = False
        else:
            self._comm.step(g.g32, g.p32, m, v, g.p16, norm, hyper)

    def reduced_gradient(self) -> torch.Tensor:
        """fp32 view of the cross-rank gradient SUM left by :meth:`backward_last` (overlapped mode):
        the communicator's symmetric buffer on the peer-to-peer path, the arena itself in-switch.
        Synchronise the device and pass a barrier before reading it."""
        lib = self._comm._lib
        lib.llmtrain_b200_nvl_reduced.restype = ctypes.c_void_p
        lib.llmtrain_b200_nvl_reduced.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        address = int(lib.llmtrain_b200_nvl_reduced(self._comm._handle, ptr(self._gpt.g32)))
        n = self._gpt.n_params

        class _Raw:
            __cuda_array_interface__ = {
                "shape": (n,), "typestr": "<f4", "data": (address, False), "version": 3,
                "strides": None,
            }

        raw = _Raw()
        tensor = torch.as_tensor(raw, device=torch.device("cuda", torch.cuda.current_device()))
        tensor._owner = (self, raw)  # type: ignore[attr-defined]
        return tensor

    def close(self) -> None:
        if self in _live:
            _live.remove(self)
        if self._team is not None:
            check(self._comm._lib.llmtrain_b200_nvl_attach_nvls(self._comm._handle, None, None),
                  "nvl_attach_nvls")
            self._gpt.g32 = None  # the arena lived in the team's buffer
            self._team.close()
            self._team = None
        self._comm.close()


def _want_nvls(world: int, overlap: bool) -> bool:
    """In-switch reduction pays from 4 GPUs up (at 2 the bf16 peer-to-peer wire format is cheaper);
    LLMTRAIN_NVLS=0/1 forces the choice."""
    import os

    forced = os.environ.get("LLMTRAIN_NVLS")
    if forced is not None:
        return overlap and forced != "0"
    return overlap and world >= 4


def try_create(gpt: Any, m: torch.Tensor, v: torch.Tensor, *, rank: int, world: int,
               device: torch.device, timeout_s: int, overlap: bool = True,
               allow_nvls: bool = True) -> Any | None:
    """Build the peer-memory communicator, or ``None`` when the ranks cannot map each other."""
    del m, v, device
    if not 2 <= world <= 8:
        return None
    ok = torch.tensor([1 if _peers_reachable(world) else 0], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok.item()) == 0:
        return None
    try:
        comm = NvlComm(gpt.n_params, rank=rank, world=world, timeout_s=timeout_s)
    except RuntimeError as exc:  # raised on every rank together (see _collective_stage)
        logger.warning("nvl communicator could not be created: %s", exc)
        return None
    team = None
    # (deterministic runs keep the peer-to-peer path: its sums are rank-ordered by construction,
    # the order of additions inside the switch is not specified)
    if allow_nvls and _want_nvls(world, overlap) and gpt.n_params % (4 * world) == 0:
        team = NvlsTeam.try_create(comm, gpt.n_params)
    return _NvlTrainComm(gpt, comm, overlap=overlap, team=team)


def shutdown() -> None:
    """Release peer mappings (called from ``teardown_ddp``)."""
    while _live:
        comm = _live.pop()
        try:
            comm.close()
        except Exception:  # pragma: no cover
            logger.debug("nvl close failed", exc_info=True)
