"""Tensor-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch).

Mirrors the slice of python/sglang/srt/distributed the hot path uses:
  tensor_model_parallel_all_reduce   communication_op.py:11-13 -> GroupCoordinator.all_reduce parallel_state.py:426-500
  get_tensor_model_parallel_{rank,world_size}                     parallel_state.py
The reference routes <8 MiB messages to a custom P2P all-reduce and the rest to NCCL
(custom_all_reduce.py:148,414-436).  Same split here: half-precision CUDA messages up to B200_AR_MAX_BYTES
(default 8 MiB) go through the one-shot NVLink peer-memory kernel of csrc/allreduce.cu (CustomAllReduce below,
one launch, graph-capturable, bit-identical results on every rank); everything else is ncclAllReduce(sum) in
place on the compute stream.  KV heads shard across ranks with no communication; only the row-parallel
o_proj / down_proj outputs are reduced.  B200_CUSTOM_AR=0 keeps every message on NCCL.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist

_group: Optional[dist.ProcessGroup] = None
_custom_ar: Optional["CustomAllReduce"] = None


class CustomAllReduce:
    """custom_all_reduce.py:CustomAllreduce for one node of NVLink-connected GPUs, one process per GPU.

    __init__ is collective: every rank allocates its region in the C library, the 64-byte CUDA-IPC handles travel
    through torch.distributed, and each rank maps its peers' regions (custom_all_reduce.py:200-260 does the same with
    its meta / buffer pointers)."""

    def __init__(self, group, device: torch.device, max_bytes: int = 8 << 20):
        import ctypes as C

        from . import _lib
        from .ops import check

        self.lib = _lib.load()
        self.device = device
        self.max_bytes = int(max_bytes)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if not (2 <= self.world <= 8):
            raise ValueError("CustomAllReduce: 2..8 ranks of one node")
        with torch.cuda.device(device):
            region = C.c_void_p()
            handle = C.create_string_buffer(64)
            check(self.lib.b200_ar_alloc(self.max_bytes, C.byref(region), handle))
            mine = torch.frombuffer(bytearray(handle.raw), dtype=torch.uint8).to(device)
            gathered = [torch.empty(64, dtype=torch.uint8, device=device) for _ in range(self.world)]
            dist.all_gather(gathered, mine, group=group)
            blob = bytes(torch.stack(gathered).cpu().numpy().tobytes())
            ctx = C.c_void_p()
            check(self.lib.b200_ar_init(region, blob, self.rank, self.world, self.max_bytes, C.byref(ctx)))
            self.ctx = ctx
            dist.barrier(group=group)   # nobody launches before every mapping exists

    def should_custom_ar(self, t: torch.Tensor) -> bool:
        """custom_all_reduce.py:414-436: dtype / size / alignment gate."""
        n = t.numel() * t.element_size()
        return (t.is_cuda and t.dtype in (torch.bfloat16, torch.float16) and t.is_contiguous() and n <= self.max_bytes
                and t.numel() % 8 == 0 and t.data_ptr() % 16 == 0 and self.lib.b200_ar_mode(self.world, n) != 0)

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        from .ops import DTYPE_CODE, check, current_stream

        with torch.cuda.device(t.device):
            check(self.lib.b200_ar_allreduce(self.ctx, t.data_ptr(), t.data_ptr(), t.numel(), DTYPE_CODE[t.dtype],
                                             current_stream(t.device)))
        return t

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
    leaves every all-reduce on NCCL -- NCCL is the reference's own path for the same op, not a CPU fallback."""
    global _custom_ar
    max_bytes = int(os.environ.get("B200_AR_MAX_BYTES", str(8 << 20)))
    ok = torch.ones(1, device=device)
    try:
        ar = CustomAllReduce(_group, device, max_bytes)
    except Exception as e:  # noqa: BLE001
        ar = None
        ok.zero_()
        print(f"[iaas_sglang_b200.tp] custom all-reduce unavailable on rank {dist.get_rank()}: {e}; using NCCL", flush=True)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=_group)   # all ranks or none
    _custom_ar = ar if ok.item() > 0 else None
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


def barrier() -> None:
    if dist.is_initialized():
        dist.barrier()


def shard_range(total: int, rank: Optional[int] = None, world: Optional[int] = None):
    """[begin, end) of a dimension sharded evenly over the TP ranks (column / row parallel, KV heads)."""
    rank = get_tensor_model_parallel_rank() if rank is None else rank
    world = get_tensor_model_parallel_world_size() if world is None else world
    if total % world != 0:
        raise ValueError(f"cannot shard {total} over {world} ranks")
    per = total // world
    return rank * per, (rank + 1) * per
