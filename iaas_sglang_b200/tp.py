"""Tensor-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch).

Mirrors the slice of python/sglang/srt/distributed the hot path uses:
  tensor_model_parallel_all_reduce   communication_op.py:11-13 -> GroupCoordinator.all_reduce parallel_state.py:426-500
  get_tensor_model_parallel_{rank,world_size}                     parallel_state.py
  fused all-reduce + residual-add + RMSNorm                       layers/flashinfer_comm_fusion.py:122, layers/communicator.py
The reference routes <8 MiB messages to a custom P2P all-reduce and the rest to NCCL
(custom_all_reduce.py:148,414-436).  Same split here: half-precision CUDA messages up to B200_AR_MAX_BYTES
(default 8 MiB) go through the kernels of csrc/allreduce.cu (CustomAllReduce below: one launch, graph-capturable,
identical results on every rank); everything else is ncclAllReduce(sum) in place on the compute stream.  KV heads
shard across ranks with no communication; only the row-parallel o_proj / down_proj outputs are reduced.
B200_CUSTOM_AR=0 keeps every message on NCCL; B200_AR_SYMM=0 skips the symmetric-memory (multicast) set-up.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import torch
import torch.distributed as dist

_group: Optional[dist.ProcessGroup] = None
_custom_ar: Optional["CustomAllReduce"] = None


def _all_ok(ok: bool, device: torch.device, group) -> bool:
    """Collective vote: True only if every rank says True.  Every stage of the set-up ends in one, so a rank that
    failed never leaves its peers inside a different collective."""
    t = torch.tensor([1.0 if ok else 0.0], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return bool(t.item() > 0)


class CustomAllReduce:
    """custom_all_reduce.py:CustomAllreduce for one node of NVLink-connected GPUs, one process per GPU.

    __init__ is collective.  Preferred set-up: ONE symmetric-memory allocation (torch.distributed._symmetric_memory:
    plumbing -- it exchanges the handles and returns every rank's mapping plus the NVSwitch MULTICAST address of the
    buffer, which the NVLS kernel needs).  Otherwise every rank cudaMallocs its region in the C library and the
    64-byte CUDA-IPC handles travel through torch.distributed (custom_all_reduce.py:200-260 does the same with its
    meta / buffer pointers).  Every stage is followed by a vote; a failure anywhere raises on EVERY rank."""

    def __init__(self, group, device: torch.device, max_bytes: int = 8 << 20):
        from . import _lib

        self.lib = _lib.load()
        self.device = device
        self.group = group
        self.max_bytes = int(max_bytes)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.ctx = None
        self.setup = None
        self._symm = None   # keeps the symmetric allocation and its handle alive
        if not (2 <= self.world <= 8):
            raise ValueError("CustomAllReduce: 2..8 ranks of one node")
        with torch.cuda.device(device):
            if os.environ.get("B200_AR_SYMM", "1") != "0":
                self._init_symmetric()
            if self.ctx is None:
                self._init_ipc()
        self.multicast = bool(self.lib.b200_ar_has_multicast(self.ctx))

    def _init_symmetric(self) -> None:
        lib, dev, ctx, err = self.lib, self.device, C.c_void_p(), None
        try:
            import torch.distributed._symmetric_memory as symm

            nbytes = int(lib.b200_ar_region_bytes(self.max_bytes))
            buf = symm.empty(nbytes, dtype=torch.uint8, device=dev)
            buf.zero_()
            torch.cuda.synchronize(dev)
            hdl = symm.rendezvous(buf, self.group)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            if os.environ.get("B200_AR_MULTICAST", "1") == "0":
                mc = 0
            arr = (C.c_uint64 * self.world)(*ptrs)
            rc = lib.b200_ar_init_ptrs(arr, C.c_uint64(mc), self.rank, self.world, self.max_bytes, C.byref(ctx))
            if rc != 0:
                raise RuntimeError(lib.b200_last_error().decode())
            self._symm = (buf, hdl)
        except Exception as e:  # noqa: BLE001  (any failure -> vote -> IPC path on every rank)
            err = e
        if _all_ok(err is None, dev, self.group):
            self.ctx, self.setup = ctx, "symmetric-memory"
            dist.barrier(group=self.group)   # every region is zeroed and mapped before anybody launches
            return
        if err is None:   # a peer failed: drop ours too
            lib.b200_ar_destroy(ctx)
            self._symm = None
        elif self.rank == 0 or os.environ.get("B200_AR_VERBOSE"):
            print(f"[iaas_sglang_b200.tp] symmetric-memory set-up unavailable on rank {self.rank}: {err}; trying CUDA IPC", flush=True)

    def _init_ipc(self) -> None:
        from ._lib import check

        lib, dev = self.lib, self.device
        region, handle, err = C.c_void_p(), C.create_string_buffer(64), None
        try:
            check(lib.b200_ar_alloc(self.max_bytes, C.byref(region), handle))
        except Exception as e:  # noqa: BLE001
            err = e
        if not _all_ok(err is None, dev, self.group):
            if err is None:
                lib.b200_ar_free_region(region)
            raise RuntimeError(f"CustomAllReduce: region allocation failed on some rank ({err})")
        mine = torch.frombuffer(bytearray(handle.raw), dtype=torch.uint8).to(dev)
        gathered = [torch.empty(64, dtype=torch.uint8, device=dev) for _ in range(self.world)]
        dist.all_gather(gathered, mine, group=self.group)
        blob = bytes(torch.stack(gathered).cpu().numpy().tobytes())
        ctx = C.c_void_p()
        try:
            check(lib.b200_ar_init(region, blob, self.rank, self.world, self.max_bytes, C.byref(ctx)))
        except Exception as e:  # noqa: BLE001
            err = e
        if not _all_ok(err is None, dev, self.group):
            if err is None:
                lib.b200_ar_destroy(ctx)       # closes the peer mappings and frees the region
            else:
                lib.b200_ar_free_region(region)
            raise RuntimeError(f"CustomAllReduce: opening the peers' regions failed on some rank ({err})")
        self.ctx, self.setup = ctx, "cuda-ipc"
        dist.barrier(group=self.group)   # nobody launches before every mapping exists

    # ------------------------------------------------------------------ dispatch
    def should_custom_ar(self, t: torch.Tensor) -> bool:
        """custom_all_reduce.py:414-436: dtype / size / alignment gate + the measured (world, size) policy."""
        n = t.numel() * t.element_size()
        return (t.is_cuda and t.dtype in (torch.bfloat16, torch.float16) and t.is_contiguous() and 0 < n <= self.max_bytes
                and t.numel() % 8 == 0 and t.data_ptr() % 16 == 0
                and self.lib.b200_ar_mode2(self.world, n, int(self.multicast)) != 0)

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        from ._lib import DTYPE_CODE, check, current_stream

        with torch.cuda.device(t.device):
            check(self.lib.b200_ar_allreduce(self.ctx, t.data_ptr(), t.data_ptr(), t.numel(), DTYPE_CODE[t.dtype],
                                             current_stream(t.device)))
        return t

    def all_reduce_add_rmsnorm(self, x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, eps: float, *,
                               quant_fp8: bool = False, want_half: bool = True, quant_eps: float = 1e-10):
        """One kernel: residual += all_reduce(x); y = RMSNorm(residual) * weight.  Returns y, or (x_q, x_s) with
        quant_fp8 (and y too with want_half), laid out as ops.rmsnorm_quant_fp8 returns them.  Raises
        NotImplementedError where the two separate kernels are the better plan (the caller falls back to them)."""
        from ._lib import DTYPE_CODE, check, current_stream, ptr
        from .ops import FP8_MAX, _fp8_group_outputs

        M, H = x.shape
        out = torch.empty_like(x) if (want_half or not quant_fp8) else None
        x_q = x_s = None
        if quant_fp8:
            x_q, x_s = _fp8_group_outputs(M, H, x.device)
        with torch.cuda.device(x.device):
            check(self.lib.b200_ar_allreduce_add_rmsnorm(
                self.ctx, ptr(x), ptr(residual), ptr(weight), ptr(out), ptr(x_q), H, ptr(x_s),
                x_s.stride(0) if x_s is not None else 0, x_s.stride(1) if x_s is not None else 0, M, H,
                DTYPE_CODE[x.dtype], eps, quant_eps, -FP8_MAX, FP8_MAX, current_stream(x.device)))
        if quant_fp8:
            return (x_q, x_s, out) if want_half else (x_q, x_s)
        return out

    def set_mode(self, mode: int) -> None:
        """Force a kernel (1 one-shot, 2 two-shot, 3 NVLS, 4 row-layout one-shot, 5 / 6 push protocol one- / two-shot;
        -1 = measured policy).  Collective
        by convention: every rank must set the same mode."""
        from ._lib import check

        check(self.lib.b200_ar_set_mode(self.ctx, int(mode)))

    def error(self) -> int:
        return int(self.lib.b200_ar_error(self.ctx))


def init_distributed(backend: Optional[str] = None, device: Optional[torch.device] = None) -> None:
    """Initialise from the torchrun environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT)."""
    global _group
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1 or dist.is_initialized():
        _group = dist.group.WORLD if dist.is_initialized() else None
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29531")
    kwargs = {}
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        kwargs["device_id"] = torch.device("cuda", local)
    dist.init_process_group(backend=backend, rank=int(os.environ.get("RANK", "0")), world_size=world, **kwargs)
    _group = dist.group.WORLD
    if backend == "nccl" and os.environ.get("B200_CUSTOM_AR", "1") != "0" and world <= 8:
        init_custom_all_reduce(torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))


def init_custom_all_reduce(device: torch.device) -> Optional[CustomAllReduce]:
    """Collective.  A failure to set the peer mappings up (no P2P between the GPUs, IPC refused) is reported once and
    leaves every all-reduce on NCCL -- NCCL is the reference's own path for the same op, not a CPU fallback.  The
    constructor votes after every stage, so it raises on all ranks or on none."""
    global _custom_ar
    max_bytes = int(os.environ.get("B200_AR_MAX_BYTES", str(8 << 20)))
    try:
        _custom_ar = CustomAllReduce(_group, device, max_bytes)
    except Exception as e:  # noqa: BLE001
        _custom_ar = None
        if dist.get_rank() == 0:
            print(f"[iaas_sglang_b200.tp] custom all-reduce unavailable: {e}; using NCCL", flush=True)
    return _custom_ar


def custom_all_reduce() -> Optional[CustomAllReduce]:
    return _custom_ar


def get_tensor_model_parallel_world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def get_tensor_model_parallel_rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """communication_op.py:11-13; in place like pynccl.all_reduce (pynccl.py:125-139)."""
    if get_tensor_model_parallel_world_size() == 1:
        return input_
    if _custom_ar is not None and _custom_ar.should_custom_ar(input_):
        return _custom_ar.all_reduce_(input_)
    dist.all_reduce(input_, op=dist.ReduceOp.SUM, group=_group)
    return input_


def tensor_model_parallel_fused_allreduce_rmsnorm(input_: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                                                  eps: float = 1e-6, *, quant_fp8: bool = False, want_half: bool = True):
    """flashinfer_allreduce_add_rmsnorm(input, residual, weight, eps) (layers/flashinfer_comm_fusion.py:122-170; used by
    layers/communicator.py / layernorm.py forward_with_allreduce_fusion): all-reduce `input_` over the TP group, add it
    to `residual` in place and return RMSNorm(residual) * weight -- one kernel where the custom all-reduce can serve the
    message, otherwise the all-reduce followed by the fused add + norm kernel (same result).  With quant_fp8 the group-128
    FP8 quantisation of the normalised row is returned as (x_q, x_s[, y]) for an FP8 block linear."""
    from . import ops

    ar = _custom_ar
    if (get_tensor_model_parallel_world_size() > 1 and ar is not None and input_.dim() == 2 and ar.should_custom_ar(input_)
            and input_.shape[1] % 128 == 0 and input_.shape[1] <= 8192):
        try:
            return ar.all_reduce_add_rmsnorm(input_, residual, weight, eps, quant_fp8=quant_fp8, want_half=want_half)
        except NotImplementedError:
            pass   # a function of (world, size) only: every rank takes the same branch
    x = tensor_model_parallel_all_reduce(input_)
    if quant_fp8:
        return ops.rmsnorm_quant_fp8(x, weight, eps, residual=residual, want_half=want_half)
    return ops.rmsnorm(x, weight, eps, residual=residual)


def check_health() -> None:
    """Raise if a peer failed to arrive in some all-reduce (the kernel gave up after ~20 s and the results since are
    wrong).  One 4-byte D2H copy: call it once per step at most (it synchronises the device)."""
    if _custom_ar is not None and _custom_ar.error() != 0:
        raise RuntimeError("iaas_sglang_b200.tp: a peer did not arrive in a custom all-reduce (b200_ar_error)")


def barrier() -> None:
    if dist.is_initialized():
        dist.barrier()


def shard_range(total: int, rank: Optional[int] = None, world: Optional[int] = None) -> Tuple[int, int]:
    """[begin, end) of a dimension sharded evenly over the TP ranks (column / row parallel, KV heads)."""
    rank = get_tensor_model_parallel_rank() if rank is None else rank
    world = get_tensor_model_parallel_world_size() if world is None else world
    if total % world != 0:
        raise ValueError(f"cannot shard {total} over {world} ranks")
    per = total // world
    return rank * per, (rank + 1) * per
