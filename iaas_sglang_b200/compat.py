"""Field-for-field stand-ins of the reference types the attention plugin reads, for boxes where
`sglang` is not installed (tests, bench).  When `sglang` IS importable the backend works on the real
objects: it only duck-types the attributes listed here.

Reference definitions:
  ForwardMode        python/sglang/srt/model_executor/forward_batch_info.py:61-125
  ForwardBatch       python/sglang/srt/model_executor/forward_batch_info.py:151-300
  ReqToTokenPool     python/sglang/srt/mem_cache/memory_pool.py:49-97
  MHATokenToKVPool   python/sglang/srt/mem_cache/memory_pool.py:176-466
  MLATokenToKVPool   python/sglang/srt/mem_cache/memory_pool.py:735-935
  RadixAttention     python/sglang/srt/layers/radix_attention.py:37-108
The pools are storage only (the scheduler/allocator side is out of scope): slot 0 is the reserved
padding slot (memory_pool.py:236-239, allocator.py:122-123).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from enum import IntEnum, auto
from typing import Any, List, Optional

import torch


class ForwardMode(IntEnum):
    EXTEND = auto()
    DECODE = auto()
    MIXED = auto()
    IDLE = auto()
    TARGET_VERIFY = auto()
    DRAFT_EXTEND = auto()
    DUMMY_FIRST = auto()

    def is_prefill(self):
        return self.is_extend()

    def is_extend(self):
        return self in (ForwardMode.EXTEND, ForwardMode.MIXED, ForwardMode.DRAFT_EXTEND, ForwardMode.TARGET_VERIFY)

    def is_decode(self):
        return self == ForwardMode.DECODE

    def is_mixed(self):
        return self == ForwardMode.MIXED

    def is_idle(self):
        return self == ForwardMode.IDLE

    def is_target_verify(self):
        return self == ForwardMode.TARGET_VERIFY

    def is_draft_extend(self):
        return self == ForwardMode.DRAFT_EXTEND

    def is_cuda_graph(self):
        return self in (ForwardMode.DECODE, ForwardMode.TARGET_VERIFY, ForwardMode.IDLE)

    def is_decode_or_idle(self):
        return self in (ForwardMode.DECODE, ForwardMode.IDLE)


class ReqToTokenPool:
    """req_to_token int32 [size, max_context_len], zero-initialised (memory_pool.py:68-70)."""

    def __init__(self, size: int, max_context_len: int, device, enable_memory_saver: bool = False):
        self.size = size
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((size, max_context_len), dtype=torch.int32, device=device)
        self.free_slots = list(range(size))

    def write(self, indices, values):
        self.req_to_token[indices] = values

    def available_size(self):
        return len(self.free_slots)

    def alloc(self, need_size: int) -> Optional[List[int]]:
        if need_size > len(self.free_slots):
            return None
        sel, self.free_slots = self.free_slots[:need_size], self.free_slots[need_size:]
        return sel

    def free(self, idx):
        self.free_slots.extend([idx] if isinstance(idx, int) else list(idx))

    def clear(self):
        self.free_slots = list(range(self.size))


class MHATokenToKVPool:
    """Per-layer k_buffer/v_buffer [size + page_size, head_num, head_dim] (memory_pool.py:236-253)."""

    def __init__(self, size: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int, layer_num: int,
                 device, enable_memory_saver: bool = False, start_layer: Optional[int] = None,
                 end_layer: Optional[int] = None, v_head_dim: Optional[int] = None):
        self.size, self.page_size, self.dtype, self.device = size, page_size, dtype, device
        self.store_dtype = dtype
        self.head_num, self.head_dim, self.layer_num = head_num, head_dim, layer_num
        self.v_head_dim = v_head_dim or head_dim
        self.start_layer = start_layer or 0
        self.end_layer = end_layer or layer_num - 1
        # memory_pool.py:116-121: fp8 pools are stored as uint8 (index_put has no fp8) and viewed as the dtype on read
        self.store_dtype = torch.uint8 if dtype in (torch.float8_e4m3fn, torch.float8_e5m2) else dtype
        self.k_buffer = [torch.zeros((size + page_size, head_num, head_dim), dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]
        self.v_buffer = [torch.zeros((size + page_size, head_num, self.v_head_dim), dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]

    def get_key_buffer(self, layer_id: int):
        b = self.k_buffer[layer_id - self.start_layer]
        return b.view(self.dtype) if self.store_dtype != self.dtype else b

    def get_value_buffer(self, layer_id: int):
        b = self.v_buffer[layer_id - self.start_layer]
        return b.view(self.dtype) if self.store_dtype != self.dtype else b

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer, loc, cache_k, cache_v, k_scale=None, v_scale=None, layer_id_override=None):
        """memory_pool.py:418-456 (same-dtype branch), on our scatter kernel."""
        from . import ops

        lid = layer_id_override if layer_id_override is not None else layer.layer_id
        if cache_k.dtype != self.dtype:   # memory_pool.py:431-440: divide by the scale, cast to the cache dtype
            if self.dtype != torch.float8_e4m3fn:
                raise NotImplementedError("KV cache dtype cast other than bf16/f16 -> float8_e4m3fn")
            ops.set_kv_buffer_fp8(self.get_key_buffer(lid), self.get_value_buffer(lid), loc, cache_k, cache_v, k_scale, v_scale)
            return
        ops.set_kv_buffer(self.get_key_buffer(lid), self.get_value_buffer(lid), loc, cache_k, cache_v)


class MLATokenToKVPool:
    """kv_buffer [size + page_size, 1, kv_lora_rank + qk_rope_head_dim] per layer (memory_pool.py:781-788);
    get_value_buffer is the [..., :kv_lora_rank] view of the same storage (:829-837)."""

    def __init__(self, size: int, page_size: int, dtype: torch.dtype, kv_lora_rank: int, qk_rope_head_dim: int,
                 layer_num: int, device, enable_memory_saver: bool = False, start_layer: Optional[int] = None,
                 end_layer: Optional[int] = None):
        self.size, self.page_size, self.dtype, self.device = size, page_size, dtype, device
        self.store_dtype = dtype
        self.kv_lora_rank, self.qk_rope_head_dim, self.layer_num = kv_lora_rank, qk_rope_head_dim, layer_num
        self.start_layer = start_layer or 0
        self.end_layer = end_layer or layer_num - 1
        self.kv_buffer = [torch.zeros((size + page_size, 1, kv_lora_rank + qk_rope_head_dim), dtype=dtype, device=device)
                          for _ in range(layer_num)]

    def get_key_buffer(self, layer_id: int):
        return self.kv_buffer[layer_id - self.start_layer]

    def get_value_buffer(self, layer_id: int):
        return self.kv_buffer[layer_id - self.start_layer][..., : self.kv_lora_rank]

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer, loc, cache_k, cache_v):
        from . import ops

        buf = self.get_key_buffer(layer.layer_id)
        ops.set_kv_buffer(buf, buf, loc, cache_k, None)

    def set_mla_kv_buffer(self, layer, loc, cache_k_nope, cache_k_rope):
        """memory_pool.py:859-876: row = concat(k_nope [.., kv_lora_rank], k_rope [.., rope_dim])."""
        B = loc.numel()
        row = torch.cat([cache_k_nope.reshape(B, 1, -1), cache_k_rope.reshape(B, 1, -1)], dim=-1)
        self.set_kv_buffer(layer, loc, row, None)


class RadixAttention:
    """The attributes of layers/radix_attention.py:37-81 the backends read."""

    def __init__(self, num_heads: int, head_dim: int, scaling: float, num_kv_heads: int, layer_id: int,
                 logit_cap: float = 0.0, v_head_dim: int = -1, sliding_window_size: int = -1,
                 is_cross_attention: bool = False):
        self.tp_q_head_num = num_heads
        self.tp_k_head_num = num_kv_heads
        self.tp_v_head_num = num_kv_heads
        self.head_dim = head_dim
        self.qk_head_dim = head_dim
        self.v_head_dim = v_head_dim if v_head_dim != -1 else head_dim
        self.scaling = scaling
        self.layer_id = layer_id
        self.logit_cap = logit_cap
        self.sliding_window_size = sliding_window_size or -1
        self.is_cross_attention = is_cross_attention
        self.k_scale = None
        self.v_scale = None

    def forward(self, q, k, v, forward_batch: "ForwardBatch", save_kv_cache: bool = True, **kwargs):
        if k is not None:  # radix_attention.py:90-98
            assert v is not None
            if "k_rope" not in kwargs:
                k = k.view(-1, self.tp_k_head_num, self.qk_head_dim)
                v = v.view(-1, self.tp_v_head_num, self.v_head_dim)
            else:
                k = k.view(-1, self.tp_k_head_num, self.v_head_dim)
        return forward_batch.attn_backend.forward(q, k, v, self, forward_batch, save_kv_cache, **kwargs)

    __call__ = forward


@dataclass
class ForwardBatch:
    """The fields of forward_batch_info.py:151-300 that an attention backend touches."""

    forward_mode: ForwardMode
    batch_size: int
    input_ids: Optional[torch.Tensor] = None
    req_pool_indices: Optional[torch.Tensor] = None  # int64 [B]
    seq_lens: Optional[torch.Tensor] = None          # int64 [B]
    out_cache_loc: Optional[torch.Tensor] = None     # int64 [B] (decode) / [T_ext] (extend)
    seq_lens_sum: int = 0
    seq_lens_cpu: Optional[torch.Tensor] = None
    positions: Optional[torch.Tensor] = None
    extend_num_tokens: Optional[int] = None
    extend_seq_lens: Optional[torch.Tensor] = None
    extend_prefix_lens: Optional[torch.Tensor] = None
    extend_start_loc: Optional[torch.Tensor] = None
    extend_prefix_lens_cpu: Optional[List[int]] = None
    extend_seq_lens_cpu: Optional[List[int]] = None
    encoder_lens: Optional[torch.Tensor] = None
    req_to_token_pool: Any = None
    token_to_kv_pool: Any = None
    attn_backend: Any = None
    spec_info: Any = None
    spec_algorithm: Any = None


@dataclass
class ModelConfigStub:
    """model_runner.model_config attributes read by the backends (triton_backend.py:41-122)."""

    num_attention_heads: int
    num_key_value_heads: int
    head_dim: int
    context_len: int
    v_head_dim: Optional[int] = None
    attention_arch_is_mla: bool = False
    kv_lora_rank: int = 0
    qk_rope_head_dim: int = 0

    def get_num_kv_heads(self, tp_size: int) -> int:
        return max(1, self.num_key_value_heads // tp_size)  # configs/model_config.py:334-341


@dataclass
class ServerArgsStub:
    attention_backend: str = "b200"
    page_size: int = 1
    triton_attention_num_kv_splits: int = 8
    disable_cuda_graph: bool = False
    enable_deterministic_inference: bool = False


@dataclass
class ModelRunnerStub:
    """MockModelRunner of python/sglang/test/attention/test_flashattn_backend.py:15-69."""

    model_config: ModelConfigStub
    req_to_token_pool: Any
    token_to_kv_pool: Any
    device: Any = "cuda"
    tp_size: int = 1
    gpu_id: int = 0
    sliding_window_size: Optional[int] = None
    server_args: ServerArgsStub = field(default_factory=ServerArgsStub)
    dtype: torch.dtype = torch.bfloat16
    kv_cache_dtype: torch.dtype = torch.bfloat16
    page_size: int = 1
    is_hybrid: bool = False
