"""B200AttnBackend -- the attention plugin for `--attention-backend b200`.

Keeps the reference's plugin surface (python/sglang/srt/layers/attention/base_attn_backend.py:14-115):
__init__(model_runner), init_forward_metadata, init_cuda_graph_state,
init_forward_metadata_{capture,replay}_cuda_graph, get_cuda_graph_seq_len_fill_value, forward,
forward_decode, forward_extend, support_triton -- same names, arguments and error behaviour as the
in-tree backends it is modelled on (intel_amx_backend.py:15-128 for the call shape,
triton_backend.py:40-733 for the GPU contracts).

Differences that matter for speed, not for results:
  * the kernels read req_to_token / req_pool_indices / seq_lens IN PLACE (like decode.cpp:936), so
    init_forward_metadata builds no kv_indptr / kv_indices copy (triton_backend.py:172-185 copies the
    whole page table every step) and CUDA-graph replay needs no metadata kernel at all;
  * KV append (set_kv_buffer) is fused into the decode kernel;
  * scratch is sized once per step (or once per graph) and reused by every layer.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch

from . import ops

try:  # subclass the real ABC when the reference is installed, else a duck type with the same surface
    from sglang.srt.layers.attention.base_attn_backend import AttentionBackend as _Base  # type: ignore
except Exception:  # noqa: BLE001
    class _Base:  # minimal copy of the dispatch in base_attn_backend.py:57-87
        def forward(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
            if forward_batch.forward_mode.is_decode():
                return self.forward_decode(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)
            return self.forward_extend(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)


@dataclass
class ForwardMetadata:
    attn_logits: Optional[torch.Tensor]  # fp32 scratch for the KV splits (intel_amx_backend.py:36-46 keeps the name)
    num_kv_splits: int
    max_seq_len: int
    max_extend_len: Optional[int]


class B200AttnBackend(_Base):
    def __init__(self, model_runner, skip_prefill: bool = False, kv_indptr_buf=None, cascade: Optional[bool] = None):
        super().__init__()
        self.forward_metadata: Optional[ForwardMetadata] = None
        self.device = model_runner.device
        mc = model_runner.model_config
        self.num_head = mc.num_attention_heads // model_runner.tp_size
        self.num_kv_head = mc.get_num_kv_heads(model_runner.tp_size)
        self.v_head_dim = model_runner.token_to_kv_pool.get_value_buffer(0).shape[-1]
        self.head_dim = model_runner.token_to_kv_pool.get_key_buffer(0).shape[-1]
        self.max_context_len = mc.context_len
        self.req_to_token = model_runner.req_to_token_pool.req_to_token
        sw = getattr(model_runner, "sliding_window_size", None)
        if sw is not None and sw > 0:
            raise NotImplementedError("b200 attention backend: sliding-window attention is not supported")
        self.static_kv_splits = 0  # 0 = heuristic (cf. SGLANG_TRITON_DECODE_ATTN_STATIC_KV_SPLITS, triton_backend.py:111-113)
        self.cuda_graph_attn_logits: Optional[torch.Tensor] = None
        self.cuda_graph_max_bs = 0
        # Cascade (shared-prefix) decode, SURVEY 8f rank 4: when every request of a decode batch starts with the same KV
        # slots (RadixAttention prefix sharing), the shared rows are attended to ONCE for the whole batch and merged by
        # LSE with each request's own suffix (sgl-kernel/csrc/attention/cascade.cu; FlashInfer's cascade wrapper).  Off
        # by default (it costs one small kernel per layer when nothing is shared): server_args.b200_cascade / B200_CASCADE=1
        # / cascade=True.  GQA/MHA with head_dim in {64, 128} only; anything else stays on the plain path.
        if cascade is None:
            cascade = bool(getattr(getattr(model_runner, "server_args", None), "b200_cascade", False)) or os.environ.get("B200_CASCADE", "0") == "1"
        self.cascade = bool(cascade) and self.head_dim == self.v_head_dim and self.head_dim in (64, 128)
        self.cascade_state: Optional[ops.CascadeState] = None
        self._dtype = model_runner.token_to_kv_pool.get_key_buffer(0).dtype

    # ------------------------------------------------------------------ metadata
    def _splits_and_scratch(self, bs: int, max_seq_len: int, scratch: Optional[torch.Tensor] = None):
        S = ops.decode_attention_num_splits(bs, self.num_head, self.num_kv_head, self.head_dim, self.v_head_dim,
                                            max_seq_len, self.static_kv_splits)
        need = ops.decode_attention_workspace_bytes(bs, self.num_head, self.v_head_dim, S)
        if scratch is None or scratch.numel() * scratch.element_size() < need:
            scratch = torch.empty((need + 3) // 4, dtype=torch.float32, device=self.device)
        return S, scratch

    def init_forward_metadata(self, forward_batch):
        """Once per step (model_runner.py:1415).  No device work: the page table is read in place."""
        if getattr(forward_batch, "spec_info", None) is not None:
            raise NotImplementedError("b200 attention backend: speculative decoding is not supported")
        bs = forward_batch.batch_size
        if forward_batch.forward_mode.is_decode_or_idle():
            cpu = getattr(forward_batch, "seq_lens_cpu", None)
            max_seq_len = int(cpu.max()) if cpu is not None and cpu.numel() > 0 else self.max_context_len
            prev = self.forward_metadata.attn_logits if self.forward_metadata is not None else None
            S, scratch = self._splits_and_scratch(bs, max_seq_len, prev)
            self.forward_metadata = ForwardMetadata(scratch, S, max_seq_len, None)
            if self.cascade and forward_batch.forward_mode.is_decode() and bs > 0:
                self._cascade_detect(bs, forward_batch.req_pool_indices, forward_batch.seq_lens)
        else:
            cpu = getattr(forward_batch, "extend_seq_lens_cpu", None)
            if cpu is not None:
                max_extend_len = int(max(cpu)) if len(cpu) > 0 else 0
            else:
                max_extend_len = int(torch.max(forward_batch.extend_seq_lens).item())
            self.forward_metadata = ForwardMetadata(None, 1, 0, max_extend_len)

    def _cascade_detect(self, bs: int, req_pool_indices, seq_lens):
        """once per step: the shared page-table prefix of the batch, left on the device (ops.CascadeState)"""
        st = self.cascade_state
        if st is None or st.max_bs < bs:
            st = self.cascade_state = ops.CascadeState(max(bs, self.cuda_graph_max_bs), self.num_head, self.v_head_dim, self._dtype, self.device)
        st.detect(self.req_to_token, req_pool_indices, seq_lens)

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int = 0, kv_indices_buf=None):
        """Fixed scratch for every captured batch size (cuda_graph_runner.py:140-187)."""
        self.cuda_graph_max_bs = max_bs
        need = 0
        for bs in range(1, max_bs + 1):
            S = ops.decode_attention_num_splits(bs, self.num_head, self.num_kv_head, self.head_dim, self.v_head_dim,
                                                self.max_context_len, self.static_kv_splits)
            need = max(need, ops.decode_attention_workspace_bytes(bs, self.num_head, self.v_head_dim, S))
        self.cuda_graph_attn_logits = torch.empty((need + 3) // 4, dtype=torch.float32, device=self.device)
        if self.cascade:   # static buffers: the captured kernels read the prefix length from them at replay time
            self.cascade_state = ops.CascadeState(max_bs, self.num_head, self.v_head_dim, self._dtype, self.device)

    def init_forward_metadata_capture_cuda_graph(self, bs: int, num_tokens: int, req_pool_indices, seq_lens, encoder_lens,
                                                 forward_mode, spec_info):
        if spec_info is not None:
            raise NotImplementedError("b200 attention backend: speculative decoding is not supported")
        if not forward_mode.is_decode_or_idle():
            raise ValueError(f"Invalid forward mode: {forward_mode=} for CUDA Graph capture.")
        assert self.cuda_graph_attn_logits is not None, "init_cuda_graph_state was not called"
        # the split count is baked into the captured launch: size it for the longest context
        S, _ = self._splits_and_scratch(bs, self.max_context_len, self.cuda_graph_attn_logits)
        self.forward_metadata = ForwardMetadata(self.cuda_graph_attn_logits, S, self.max_context_len, None)
        if self.cascade and forward_mode.is_decode():
            self._cascade_detect(bs, req_pool_indices, seq_lens)

    def init_forward_metadata_replay_cuda_graph(self, bs: int, req_pool_indices, seq_lens, seq_lens_sum: int, encoder_lens,
                                                forward_mode, spec_info, seq_lens_cpu=None):
        """Nothing to do: the captured kernels read req_pool_indices / seq_lens / req_to_token in place
        (the Triton backend re-runs its kv_indices builder here, triton_backend.py:532-627)."""
        if spec_info is not None:
            raise NotImplementedError("b200 attention backend: speculative decoding is not supported")
        if not forward_mode.is_decode_or_idle():
            raise ValueError("Invalid forward mode")
        if self.cascade and forward_mode.is_decode():   # eager, before graph.replay(): one scan of the page table
            self._cascade_detect(bs, req_pool_indices[:bs], seq_lens[:bs])

    def get_cuda_graph_seq_len_fill_value(self):
        return 1  # triton_backend.py:629-630

    # ------------------------------------------------------------------ forward
    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True,
                       q_rope: Optional[torch.Tensor] = None, k_rope: Optional[torch.Tensor] = None,
                       quant_fp8: bool = False):
        """base_attn_backend.py:89-99.  q_rope / k_rope are the MLA arguments of cutlass_mla_backend.py:226-285 /
        flashinfer_mla_backend.py: q then holds the 512 latent columns and q_rope the 64 rope columns of every head, k
        the latent part and k_rope the rope part of the new cache row (set_mla_kv_buffer, memory_pool.py:859-876); the
        kernel takes the two pointers of each, nothing is concatenated.  seq_lens / req_pool_indices may be int32
        (cuda_graph_runner.py:273-276) or int64.  Any other keyword is a TypeError, as in the reference backends.
        quant_fp8=True (not a reference argument): the caller is an FP8 block o_proj and wants the group-128 quantisation
        of `o` from the same launch sequence -> returns (x_q, x_s) for LinearMethod.apply_quantized."""
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("b200 attention backend: cross attention is not supported")
        md = self.forward_metadata
        H = layer.tp_q_head_num
        pool = forward_batch.token_to_kv_pool
        k_buf = pool.get_key_buffer(layer.layer_id)
        v_buf = pool.get_value_buffer(layer.layer_id)
        if q_rope is not None:
            qv = q.view(-1, H, layer.v_head_dim)
            q_rope = q_rope.view(-1, H, layer.qk_head_dim - layer.v_head_dim)
        else:
            qv = q.reshape(-1, H, layer.qk_head_dim)
        o = qv.new_empty((qv.shape[0], H * layer.v_head_dim))
        key = value = kr = None
        if save_kv_cache and k is not None:
            if k_rope is not None:
                key = k.view(-1, layer.tp_k_head_num, layer.v_head_dim)
                kr = k_rope.view(-1, layer.tp_k_head_num, layer.qk_head_dim - layer.v_head_dim)
            else:
                key = k.view(-1, layer.tp_k_head_num, layer.qk_head_dim)
                value = v.view(-1, layer.tp_v_head_num, layer.v_head_dim) if v is not None else None
        casc = None
        if (self.cascade and self.cascade_state is not None and q_rope is None and qv.shape[0] <= self.cascade_state.max_bs
                and k_buf.element_size() == 2):
            casc = self.cascade_state
            casc.prefix_attention(qv, k_buf, v_buf, forward_batch.req_to_token_pool.req_to_token, layer.scaling, layer.logit_cap)
        res = ops.decode_attention(
            qv, k_buf, v_buf, o.view(-1, H, layer.v_head_dim), key, value, forward_batch.out_cache_loc,
            md.attn_logits, forward_batch.req_to_token_pool.req_to_token, forward_batch.req_pool_indices,
            forward_batch.seq_lens, layer.scaling, layer.logit_cap,
            max_seq_len=md.max_seq_len, num_splits=md.num_kv_splits, fused_kv_write=True, quant_fp8=bool(quant_fp8),
            q_rope=q_rope, k_rope=kr, cascade=casc,
            k_scale=getattr(layer, "k_scale_float", None) or (float(layer.k_scale) if getattr(layer, "k_scale", None) is not None else None),
            v_scale=getattr(layer, "v_scale_float", None) or (float(layer.v_scale) if getattr(layer, "v_scale", None) is not None else None))
        if quant_fp8:
            return res[1], res[2]
        return o

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True,
                       q_rope: Optional[torch.Tensor] = None, k_rope: Optional[torch.Tensor] = None):
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("b200 attention backend: cross attention is not supported")
        if q_rope is not None:   # MLA prefill (not the tuned path): one packed 576-column q / k, as the pool stores them
            H = layer.tp_q_head_num
            q = torch.cat([q.view(-1, H, layer.v_head_dim), q_rope.view(-1, H, layer.qk_head_dim - layer.v_head_dim)], dim=-1)
            q = q.view(-1, H * layer.qk_head_dim)
        if k_rope is not None:
            k = torch.cat([k.view(-1, layer.tp_k_head_num, layer.v_head_dim),
                           k_rope.view(-1, layer.tp_k_head_num, layer.qk_head_dim - layer.v_head_dim)], dim=-1)
            v = k[..., : layer.v_head_dim]
        if layer.qk_head_dim != layer.v_head_dim:
            o = q.new_empty((q.shape[0], layer.tp_q_head_num * layer.v_head_dim))
        else:
            o = torch.empty_like(q)
        pool = forward_batch.token_to_kv_pool
        if save_kv_cache:
            pool.set_kv_buffer(layer, forward_batch.out_cache_loc, k, v)
        md = self.forward_metadata
        ops.extend_attention(
            q.view(-1, layer.tp_q_head_num, layer.qk_head_dim),
            k.view(-1, layer.tp_k_head_num, layer.qk_head_dim),
            v.view(-1, layer.tp_v_head_num, layer.v_head_dim),
            o.view(-1, layer.tp_q_head_num, layer.v_head_dim),
            pool.get_key_buffer(layer.layer_id), pool.get_value_buffer(layer.layer_id),
            forward_batch.req_to_token_pool.req_to_token, forward_batch.req_pool_indices, forward_batch.seq_lens,
            forward_batch.extend_seq_lens, forward_batch.extend_start_loc, md.max_extend_len, layer.scaling,
            layer.logit_cap)
        return o

    def support_triton(self):
        return False
