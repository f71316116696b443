"""Tensor-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch).

Mirrors the slice of python/sglang/srt/distributed the hot path uses:
  tensor_model_parallel_all_reduce   communication_op.py:11-13 -> GroupCoordinator.all_reduce parallel_state.py:426-500
  get_tensor_model_parallel_{rank,world_size}                     parallel_state.py
The reference routes <8 MiB messages to a custom P2P all-reduce and the rest to NCCL
(custom_all_reduce.py:148,414-436); here every all-reduce is ncclAllReduce(sum) in place on the
compute stream (graph-capturable).  KV heads shard across ranks with no communication; only the
row-parallel o_proj / down_proj outputs are reduced.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist

_group: Optional[dist.ProcessGroup] = None


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


def get_tensor_model_parallel_world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def get_tensor_model_parallel_rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """communication_op.py:11-13; in place like pynccl.all_reduce (pynccl.py:125-139)."""
    if get_tensor_model_parallel_world_size() == 1:
        return input_
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
