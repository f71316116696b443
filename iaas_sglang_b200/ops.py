"""Kernel ops with the reference's sgl_kernel signatures, backed by the sm_100a C-ABI library.

Mirrors sgl-kernel/python/sgl_kernel/gemm.py:7-31 (awq_dequantize, int8_scaled_mm,
fp8_blockwise_scaled_mm), python/sglang/srt/layers/quantization/fp8_kernel.py:273-326
(sglang_per_token_group_quant_fp8), int8_kernel.py:51-81 (per_token_quant_int8) and
python/sglang/srt/layers/attention/utils.py (kv index builders): same positional arguments, same
returned shapes/dtypes/strides, same precondition errors.  Each op allocates its output with
torch.empty and enqueues the kernel on the current CUDA stream; nothing synchronises.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import DTYPE_CODE, c_float, c_int, c_int64, c_size_t, check, current_stream, load, ptr, require_cuda

fp8_dtype = torch.float8_e4m3fn
FP8_MAX = 448.0

# ---------------------------------------------------------------------------------------------
# scratch: one zero-initialised split-K workspace per (device, stream) -- GEMMs that run concurrently on two streams
# must not share one (partial tiles and tile counters live there, b200_sgl.h); the counters are self-resetting
# ---------------------------------------------------------------------------------------------
_gemm_ws: Dict[Tuple[torch.device, int], torch.Tensor] = {}


def gemm_workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    key = (device, int(torch.cuda.current_stream(device).cuda_stream))
    ws = _gemm_ws.get(key)
    if ws is None or ws.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            # a capture stream that never ran eagerly: share a workspace of this device (graphs of one device replay in
            # stream order unless the caller overlaps them -- then reserve one per stream before capturing)
            for (d, _), w in _gemm_ws.items():
                if d == device and w.numel() >= nbytes:
                    return w
            raise RuntimeError("iaas_sglang_b200: GEMM workspace must be sized before CUDA-graph capture "
                               "(call ops.reserve_gemm_workspace)")
        ws = torch.zeros(max(nbytes, 32 << 20), dtype=torch.uint8, device=device)
        _gemm_ws[key] = ws
    return ws


def reserve_gemm_workspace(device, nbytes: int) -> None:
    gemm_workspace(torch.device(device), nbytes)


# ---------------------------------------------------------------------------------------------
# quantised GEMMs
# ---------------------------------------------------------------------------------------------
def fp8_blockwise_scaled_mm(mat_a: torch.Tensor, mat_b: torch.Tensor, scales_a: torch.Tensor, scales_b: torch.Tensor,
                            out_dtype: torch.dtype, *, rope: Optional[tuple] = None) -> torch.Tensor:
    """sgl_kernel.fp8_blockwise_scaled_mm (gemm.py:24-31; fp8_blockwise_gemm_kernel.cu:345-423).

    mat_a [M,K] e4m3 row-major; mat_b [K,N] e4m3 column-major; scales_a [M,K/128] f32 M-major;
    scales_b [K/128,N/128] f32 K-major; returns [M,N] out_dtype.  Checks follow :351-383.
    rope = (positions int64 [M], rope_cols, head_dim, theta, cos_sin): a qkv projection whose first rope_cols output
    columns (q then k heads) get the neox rotary embedding (models/llama.py:186-188 qkv_proj -> rotary_emb) -- in the
    GEMM's epilogue where the decode-shape kernel serves the shape (head_dim 128; cos_sin = rope_cos_sin(positions, ...),
    computed once per step), otherwise (or with cos_sin None) by rope_neox_ on the result: the same bits."""
    dev = require_cuda(mat_a, mat_b, scales_a, scales_b)
    if mat_a.dim() != 2 or mat_b.dim() != 2:
        raise RuntimeError("mat_a and mat_b must be a matrix")
    if mat_a.stride(1) != 1:
        raise RuntimeError("mat_a must be a row major tensor")
    if mat_b.stride(0) != 1:
        raise RuntimeError("mat_b must be a column major tensor")
    if mat_a.shape[1] != mat_b.shape[0]:
        raise RuntimeError("mat_a and mat_b shapes cannot be multiplied")
    if mat_a.dtype != fp8_dtype or mat_b.dtype != fp8_dtype:
        raise RuntimeError("mat_a and mat_b must be Float8_e4m3fn")
    if out_dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError("out_dtype must be Half or BFloat16")
    M, K = mat_a.shape
    N = mat_b.shape[1]
    kb, nb = (K + 127) // 128, (N + 127) // 128
    if scales_a.dtype != torch.float32 or scales_b.dtype != torch.float32:
        raise RuntimeError("scales_a and scales_b must be Float32")
    if tuple(scales_a.shape) != (M, kb):
        raise RuntimeError(f"scales_a must be [M, K/128] = {(M, kb)}, got {tuple(scales_a.shape)}")
    if tuple(scales_b.shape) != (kb, nb):
        raise RuntimeError(f"scales_b must be [K/128, N/128] = {(kb, nb)}, got {tuple(scales_b.shape)}")
    if M > 1 and kb > 1 and scales_a.stride(0) != 1:
        raise RuntimeError("scales_a must be M major")
    if kb > 1 and nb > 1 and scales_b.stride(0) != 1:
        raise RuntimeError("scales_b must be K major")
    out = torch.empty((M, N), dtype=out_dtype, device=dev)
    if M == 0:
        return out
    lib = load()
    nbytes = lib.b200_gemm_workspace_size(M, N, K)
    ws = gemm_workspace(dev, nbytes)
    with torch.cuda.device(dev):
        if rope is not None:
            positions, rope_cols, head_dim, theta, cos_sin = rope
            if positions.dtype != torch.int64 or not positions.is_contiguous() or positions.numel() != M or not positions.is_cuda:
                raise RuntimeError("rope positions must be a contiguous int64 CUDA tensor with M entries")
            if rope_cols % head_dim != 0 or not 0 <= rope_cols <= N:
                raise RuntimeError("rope_cols must be a multiple of head_dim, at most N")
            if cos_sin is not None:
                if (cos_sin.dtype != torch.float32 or not cos_sin.is_contiguous() or not cos_sin.is_cuda
                        or tuple(cos_sin.shape) != (M, head_dim // 2, 2)):
                    raise RuntimeError("cos_sin must be a contiguous float32 CUDA tensor [M, head_dim / 2, 2] (ops.rope_cos_sin)")
                rc = lib.b200_fp8_blockwise_scaled_mm_rope(
                    ptr(mat_a), mat_a.stride(0), ptr(mat_b), mat_b.stride(1), ptr(scales_a), scales_a.stride(0),
                    scales_a.stride(1), ptr(scales_b), scales_b.stride(0), scales_b.stride(1), ptr(out), out.stride(0),
                    DTYPE_CODE[out_dtype], M, N, K, ptr(cos_sin), int(rope_cols), int(head_dim), ptr(ws), ws.numel(),
                    current_stream(dev))
                if rc != _lib.B200_ERR_UNSUPPORTED:
                    check(rc)
                    return out
        check(lib.b200_fp8_blockwise_scaled_mm(
            ptr(mat_a), mat_a.stride(0), ptr(mat_b), mat_b.stride(1), ptr(scales_a), scales_a.stride(0),
            scales_a.stride(1), ptr(scales_b), scales_b.stride(0), scales_b.stride(1), ptr(out), out.stride(0),
            DTYPE_CODE[out_dtype], M, N, K, ptr(ws), ws.numel(), current_stream(dev)))
    if rope is not None:   # shape outside the fused epilogue: the same result from two launches
        positions, rope_cols, head_dim, theta, _ = rope
        if rope_cols > head_dim:   # (q heads and k heads are adjacent column blocks: any split of them gives the same bits)
            rope_neox_(out[:, :rope_cols - head_dim], out[:, rope_cols - head_dim:rope_cols], positions, head_dim, theta)
        elif rope_cols == head_dim:
            rope_neox_(out[:, :head_dim], out[:, :0], positions, head_dim, theta)
    return out


def rope_cos_sin(positions: torch.Tensor, head_dim: int, theta: float = 10000.0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[M, head_dim / 2, 2] float32 (cos, sin) rows of this step's positions -- rotary_embedding.py _compute_cos_sin_cache
    indexed by `positions`, in the precision rope_neox_ evaluates them.  Once per decode step; every layer's qkv GEMM
    rotates q and k with it in its epilogue (fp8_blockwise_scaled_mm(rope=...))."""
    dev = require_cuda(positions)
    if positions.dtype != torch.int64 or not positions.is_contiguous():
        raise RuntimeError("positions must be a contiguous int64 tensor")
    M = positions.numel()
    if out is None:
        out = torch.empty(M, head_dim // 2, 2, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(load().b200_rope_cos_sin(ptr(positions), M, int(head_dim), float(theta), ptr(out), current_stream(dev)))
    return out


def int8_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype, bias=None) -> torch.Tensor:
    """sgl_kernel.int8_scaled_mm (gemm.py:13-21; int8_gemm_kernel.cu:663-747). The reference has no
    sm_100 path for this op (:724-743 raises on B200); this one runs on tcgen05 kind::i8."""
    dev = require_cuda(mat_a, mat_b, scales_a, scales_b, bias)
    if mat_a.dim() != 2 or mat_b.dim() != 2:
        raise RuntimeError("mat_a and mat_b must be a matrix")
    if mat_a.stride(1) != 1:
        raise RuntimeError("mat_a must be a row major tensor")
    if mat_b.stride(0) != 1:
        raise RuntimeError("mat_b must be a column major tensor")
    if mat_a.shape[1] != mat_b.shape[0]:
        raise RuntimeError("mat_a and mat_b shapes cannot be multiplied")
    if mat_a.dtype != torch.int8 or mat_b.dtype != torch.int8:
        raise RuntimeError("mat_a and mat_b must be Int8")
    if out_dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError("out_dtype must be Half or BFloat16")
    M, K = mat_a.shape
    N = mat_b.shape[1]
    if scales_a.numel() != M or scales_b.numel() != N:
        raise RuntimeError("size of scales_a / scales_b is not matched")
    if not scales_a.is_contiguous() or not scales_b.is_contiguous():
        raise RuntimeError("scales_a and scales_b must be contiguous")
    if scales_a.dtype != torch.float32 or scales_b.dtype != torch.float32:
        raise RuntimeError("scales_a and scales_b must be Float32")
    if bias is not None:
        if bias.numel() != N or not bias.is_contiguous() or bias.dtype != out_dtype:
            raise RuntimeError("bias must be a contiguous [N] tensor of out_dtype")
    out = torch.empty((M, N), dtype=out_dtype, device=dev)
    if M == 0:
        return out
    lib = load()
    ws = gemm_workspace(dev, lib.b200_gemm_workspace_size(M, N, K))
    with torch.cuda.device(dev):
        check(lib.b200_int8_scaled_mm(ptr(mat_a), mat_a.stride(0), ptr(mat_b), mat_b.stride(1), ptr(scales_a),
                                      ptr(scales_b), ptr(bias), ptr(out), out.stride(0), DTYPE_CODE[out_dtype], M, N, K,
                                      ptr(ws), ws.numel(), current_stream(dev)))
    return out


def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """sgl_kernel.awq_dequantize (gemm.py:7-10; awq_kernel.cu:186-221) -> [rows, cols*8] of scales.dtype."""
    dev = require_cuda(qweight, scales, qzeros)
    if qweight.dtype != torch.int32 or qzeros.dtype != torch.int32:
        raise RuntimeError("qweight and qzeros must be Int32")
    if scales.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError("scales must be Half or BFloat16")
    rows, cols = qweight.shape
    group = rows // scales.shape[0]
    qweight, scales, qzeros = qweight.contiguous(), scales.contiguous(), qzeros.contiguous()
    out = torch.empty((rows, cols * 8), dtype=scales.dtype, device=dev)
    with torch.cuda.device(dev):
        check(load().b200_awq_dequantize(ptr(qweight), ptr(scales), ptr(qzeros), ptr(out), DTYPE_CODE[scales.dtype], rows,
                                         cols, group, current_stream(dev)))
    return out


_awq_ws: Dict[torch.device, torch.Tensor] = {}


def awq_w4a16_gemm(x: torch.Tensor, qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor,
                   bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """AWQLinearMethod.apply's math (awq.py:186-204): x @ dequant(qweight) (+ bias) on our own kernels."""
    dev = require_cuda(x, qweight, scales, qzeros, bias)
    if x.dim() != 2 or x.stride(1) != 1:
        raise RuntimeError("x must be a row major matrix")
    if x.dtype != scales.dtype or x.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError("x and scales must share a Half/BFloat16 dtype")
    M, K = x.shape
    if qweight.shape[0] != K:
        raise RuntimeError("x and qweight shapes cannot be multiplied")
    N = qweight.shape[1] * 8
    group = K // scales.shape[0]
    qweight, scales, qzeros = qweight.contiguous(), scales.contiguous(), qzeros.contiguous()
    out = torch.empty((M, N), dtype=x.dtype, device=dev)
    if M == 0:
        return out
    lib = load()
    need = lib.b200_awq_w4a16_workspace_size(M, N, K)
    ws = _awq_ws.get(dev)
    if ws is None or ws.numel() < need:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("iaas_sglang_b200: AWQ workspace must be sized before CUDA-graph capture")
        ws = torch.zeros(need, dtype=torch.uint8, device=dev)
        _awq_ws[dev] = ws
    with torch.cuda.device(dev):
        check(lib.b200_awq_w4a16_mm(ptr(x), x.stride(0), ptr(qweight), ptr(scales), ptr(qzeros), ptr(bias), ptr(out),
                                    out.stride(0), DTYPE_CODE[x.dtype], M, N, K, group, ptr(ws), ws.numel(),
                                    current_stream(dev)))
    return out


# ---------------------------------------------------------------------------------------------
# activation quantisation
# ---------------------------------------------------------------------------------------------
def sgl_per_token_group_quant_fp8(input: torch.Tensor, output_q: torch.Tensor, output_s: torch.Tensor, group_size: int,
                                  eps: float, fp8_min: float, fp8_max: float, scale_ue8m0: bool = False) -> None:
    """torch.ops.sgl_kernel.sgl_per_token_group_quant_fp8 (common_extension.cc:116-119): in-place form."""
    dev = require_cuda(input, output_q, output_s)
    if scale_ue8m0:
        raise NotImplementedError("scale_ue8m0 (DeepGEMM packed scales) is outside the hot path (SURVEY 2.1 #19)")
    if input.stride(-1) != 1 or output_q.stride(-1) != 1:
        raise RuntimeError("input and output_q must be contiguous in the last dimension")
    x2 = input.reshape(-1, input.shape[-1])
    q2 = output_q.view(-1, output_q.shape[-1])
    s2 = output_s.view(-1, output_s.shape[-1]) if output_s.dim() != 2 else output_s
    M, K = x2.shape
    with torch.cuda.device(dev):
        check(load().b200_per_token_group_quant_fp8(ptr(x2), x2.stride(0), DTYPE_CODE[x2.dtype], ptr(q2), q2.stride(0),
                                                    ptr(s2), s2.stride(0), s2.stride(1), M, K, group_size, eps, fp8_min,
                                                    fp8_max, current_stream(dev)))


def per_token_group_quant_fp8(x: torch.Tensor, group_size: int = 128, eps: float = 1e-10,
                              column_major_scales: bool = False, scale_tma_aligned: bool = False,
                              scale_ue8m0: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    """sglang_per_token_group_quant_fp8 (fp8_kernel.py:273-326) -> (x_q e4m3, x_s f32 [.., K/group])."""
    assert x.shape[-1] % group_size == 0, "the last dimension of `x` cannot be divisible by `group_size`"
    assert x.is_contiguous(), "`x` is not contiguous"
    if scale_ue8m0:
        raise NotImplementedError("scale_ue8m0 is outside the hot path")
    x_q = torch.empty_like(x, dtype=fp8_dtype)
    M = x.numel() // x.shape[-1]
    G = x.shape[-1] // group_size
    if column_major_scales:
        if scale_tma_aligned:
            aligned = (M + 3) // 4 * 4
            x_s = torch.empty((G, aligned), dtype=torch.float32, device=x.device).t()[:M, :]
        else:
            x_s = torch.empty((G, M), dtype=torch.float32, device=x.device).t()
    else:
        x_s = torch.empty((M, G), dtype=torch.float32, device=x.device)
    if x.numel() > 0:
        sgl_per_token_group_quant_fp8(x, x_q, x_s, group_size, eps, -FP8_MAX, FP8_MAX, False)
    if x.dim() != 2:
        x_s = x_s.reshape(x.shape[:-1] + (G,)) if not column_major_scales else x_s
    return x_q, x_s


def per_token_quant_int8(x: torch.Tensor, scale_dtype=torch.float32, cal_sum: bool = False):
    """per_token_quant_int8 (int8_kernel.py:51-81) -> (x_q int8, scales [..,1] f32)."""
    dev = require_cuda(x)
    assert x.is_contiguous()
    if cal_sum:
        raise NotImplementedError("cal_sum is only used by the w8a8-fp8/int8 zero-point paths (out of scope)")
    if scale_dtype != torch.float32:
        raise NotImplementedError("per_token_quant_int8: only float32 scales")
    x2 = x.view(-1, x.shape[-1])
    M, K = x2.shape
    x_q = torch.empty_like(x, dtype=torch.int8)
    scales = torch.empty(x.shape[:-1] + (1,), device=dev, dtype=torch.float32)
    if M > 0:
        with torch.cuda.device(dev):
            check(load().b200_per_token_quant_int8(ptr(x2), x2.stride(0), DTYPE_CODE[x.dtype], ptr(x_q), K, ptr(scales), M,
                                                   K, current_stream(dev)))
    return x_q, scales


# ---------------------------------------------------------------------------------------------
# page-table index builders
# ---------------------------------------------------------------------------------------------
def _i64(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if t is None:
        return None
    return t if t.dtype == torch.int64 and t.is_contiguous() else t.to(torch.int64).contiguous()


def create_kv_indices(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, page_kernel_lens: torch.Tensor,
                      kv_indptr: torch.Tensor, kv_start_idx: Optional[torch.Tensor], kv_indices: torch.Tensor) -> None:
    """create_flashinfer_kv_indices_triton[(bs,)](...) (attention/utils.py:5-41); writes kv_indices in place."""
    dev = require_cuda(req_to_token, req_pool_indices, page_kernel_lens, kv_indptr, kv_indices)
    if kv_indices.dtype != torch.int32 or req_to_token.dtype not in (torch.int32, torch.int64):
        raise RuntimeError("kv_indices must be int32 and req_to_token int32/int64")
    if req_to_token.stride(1) != 1:
        raise RuntimeError("req_to_token rows must be contiguous")
    rpi, lens, ip, ks = _i64(req_pool_indices), _i64(page_kernel_lens), _i64(kv_indptr), _i64(kv_start_idx)
    with torch.cuda.device(dev):
        check(load().b200_create_kv_indices(ptr(req_to_token), int(req_to_token.dtype == torch.int64),
                                            req_to_token.stride(0), ptr(rpi), ptr(lens), ptr(ip), ptr(ks), ptr(kv_indices),
                                            rpi.numel(), current_stream(dev)))


def create_page_table(req_to_token, req_pool_indices, page_kernel_lens, kv_start_idx, kv_indices: torch.Tensor,
                      page_size: int) -> None:
    """create_flashmla_kv_indices_triton[(bs,)](...) (attention/utils.py:44-94); kv_indices [B, max_pages] int32."""
    dev = require_cuda(req_to_token, req_pool_indices, page_kernel_lens, kv_indices)
    rpi, lens, ks = _i64(req_pool_indices), _i64(page_kernel_lens), _i64(kv_start_idx)
    with torch.cuda.device(dev):
        check(load().b200_create_page_table(ptr(req_to_token), int(req_to_token.dtype == torch.int64),
                                            req_to_token.stride(0), ptr(rpi), ptr(lens), ptr(ks), ptr(kv_indices),
                                            kv_indices.stride(0), rpi.numel(), page_size, kv_indices.shape[1],
                                            current_stream(dev)))


# ---------------------------------------------------------------------------------------------
# attention (functional forms; the plugin class is backend.B200AttnBackend)
# ---------------------------------------------------------------------------------------------
def decode_attention_num_splits(B: int, Hq: int, Hkv: int, D: int, Dv: int, max_seq_len: int, hint: int = 0) -> int:
    return load().b200_decode_attn_num_splits(B, Hq, Hkv, D, Dv, max_seq_len, hint)


def decode_attention_workspace_bytes(B: int, Hq: int, Dv: int, num_splits: int) -> int:
    return load().b200_decode_attn_workspace_size(B, Hq, Dv, num_splits)


_INT_META = (torch.int64, torch.int32)


def decode_attention(query, k_cache, v_cache, output, key, value, loc, attn_logits, req_to_token, req_pool_indices,
                     seq_lens, sm_scale: float, logit_cap: float, *, max_seq_len: int = 0, num_splits: int = 0,
                     fused_kv_write: bool = False, quant_fp8: bool = False, quant_eps: float = 1e-10,
                     q_rope: Optional[torch.Tensor] = None, k_rope: Optional[torch.Tensor] = None, page_size: int = 1,
                     cascade: Optional["CascadeState"] = None, k_scale: Optional[float] = None,
                     v_scale: Optional[float] = None):
    """Same positional signature as torch.ops.sgl_kernel.decode_attention_cpu (decode.cpp:1375-1388;
    schema torch_extension_cpu.cpp:264-268) on CUDA tensors.

    query [B,Hq,D]; k_cache/v_cache [T,Hkv,D]/[T,Hkv,Dv] (last dim contiguous, :1394-1399); output [B,Hq,Dv];
    key/value [B,Hkv,D]/[B,Hkv,Dv] are written to k_cache[loc]/v_cache[loc] first (None = no write);
    attn_logits = fp32 scratch (any shape, only its byte size matters; None = allocate).
    MLA is detected like the reference: k_cache and v_cache share storage, Hkv == 1, D == Dv + 64 (:1455).
    fused_kv_write=True moves the append into the attention kernel; it requires that slot loc[b] is referenced
    by request b only (true for SGLang's allocators; schedule_batch.py:1556-1568).
    seq_lens / req_pool_indices / loc are int64 as decode.cpp:1446-1450 requires, or int32 as the CUDA-graph runner
    allocates them (cuda_graph_runner.py:273-276); req_pool_indices=None means request b owns row b of the table.
    MLA extras (cutlass_mla_backend.py:226-285): q_rope [B,Hq,64] with query = q_nope [B,Hq,512]; k_rope [B,1,64] with
    key = the latent part [B,1,512] (set_mla_kv_buffer); page_size P > 1: req_to_token is a [B, pages] page table.
    FP8 KV cache (memory_pool.py:116-121,431-440): k_cache / v_cache of dtype float8_e4m3fn (or the pool's uint8 storage
    with k_scale / v_scale given) hold e4m3 codes, K = code * k_scale, V = code * v_scale; key / value stay in query's dtype
    and are appended as e4m3(x / scale).  Tuned GQA kernel, head_dim 128 (NotImplementedError otherwise).
    cascade: a CascadeState (shared_prefix_len + the prefix state of this layer): the kernel attends to the positions
    after the shared prefix only and its merge adds the prefix state -- same result, the shared KV rows are read once.
    quant_fp8=True also returns (x_q, x_s) = per_token_group_quant_fp8(output.view(B, Hq*Dv), 128,
    column_major_scales=True) for the FP8 block o_proj (fp8_utils.py:202-204), bit-identical to the separate call; the
    split-merge kernel emits it where the tuned kernel runs (Dv == 128), otherwise the quantiser is launched here.
    Returns output, or (output, x_q, x_s)."""
    dev = require_cuda(query, k_cache, v_cache, output, key, value, loc, attn_logits, req_to_token, req_pool_indices,
                       seq_lens)
    if query.dim() != 3 or k_cache.dim() != 3 or v_cache.dim() != 3:
        raise RuntimeError("query, k_cache, v_cache must be 3-d")
    if query.stride(-1) != 1 or k_cache.stride(-1) != 1 or v_cache.stride(-1) != 1:
        raise RuntimeError("query, k_cache and v_cache must be contiguous in the last dimension")
    B, Hq, D = query.shape
    Hkv, Dv = k_cache.shape[1], v_cache.shape[2]
    if q_rope is not None:
        if q_rope.dim() != 3 or q_rope.shape[:2] != query.shape[:2] or q_rope.stride(-1) != 1 or q_rope.dtype != query.dtype:
            raise RuntimeError("q_rope must be [B, Hq, rope_dim] of query's dtype, contiguous in the last dimension")
        D = D + q_rope.shape[2]
    if k_cache.shape[2] != D or v_cache.shape[1] != Hkv:
        raise RuntimeError("k_cache / v_cache shapes do not match query")
    if tuple(output.shape) != (B, Hq, Dv) or not output.is_contiguous():
        raise RuntimeError("output must be a contiguous [B, Hq, Dv] tensor")
    kv_fp8 = k_cache.dtype == torch.float8_e4m3fn or (k_cache.dtype == torch.uint8 and k_scale is not None)
    if kv_fp8:
        if v_cache.dtype != k_cache.dtype or query.dtype != output.dtype or query.dtype not in DTYPE_CODE or query.dtype == torch.float32:
            raise RuntimeError("fp8 KV cache: k_cache / v_cache must both be e4m3, query / output bf16 or f16")
    elif not (query.dtype == k_cache.dtype == v_cache.dtype == output.dtype) or query.dtype not in DTYPE_CODE:
        raise RuntimeError("query, k_cache, v_cache, output must share a bf16/f16 dtype")
    if req_to_token.dtype not in (torch.int32, torch.int64) or req_to_token.stride(1) != 1:
        raise RuntimeError("req_to_token must be int32/int64 with contiguous rows")  # decode.cpp:1441-1445
    for name, t in (("seq_lens", seq_lens), ("req_pool_indices", req_pool_indices)):
        if t is None and name == "req_pool_indices":
            continue
        if t.dtype not in _INT_META or not t.is_contiguous():
            raise RuntimeError(f"{name} must be a contiguous int64 (or int32) tensor")  # decode.cpp:1446-1450
        if t.numel() != B:
            raise RuntimeError("seq_lens / req_pool_indices must have B entries")
    write_kv = 0
    if key is not None:
        if loc is None or loc.dtype not in _INT_META or not loc.is_contiguous() or loc.numel() != B:
            raise RuntimeError("loc must be a contiguous int64 (or int32) [B] tensor")
        if key.stride(-1) != 1 or (value is not None and value.stride(-1) != 1):
            raise RuntimeError("key and value must be contiguous in the last dimension")
        if k_rope is not None and (k_rope.stride(-1) != 1 or k_rope.numel() != B * k_rope.shape[-1]
                                   or key.shape[-1] + k_rope.shape[-1] != D):
            raise RuntimeError("k_rope must be [B, 1, rope_dim] with key holding the other D - rope_dim columns")
        write_kv = 2 if fused_kv_write else 1
    lib = load()
    p = _lib.DecodeAttnParams()
    p.q, p.q_stride_b, p.q_stride_h = query.data_ptr(), query.stride(0), query.stride(1)
    if key is not None:
        p.k_new, p.knew_stride_b, p.knew_stride_h = key.data_ptr(), key.stride(0), key.stride(1)
        v_src = value if value is not None else key
        p.v_new, p.vnew_stride_b, p.vnew_stride_h = v_src.data_ptr(), v_src.stride(0), v_src.stride(1)
        p.out_cache_loc = loc.data_ptr()
    p.k_buf, p.kbuf_stride_t, p.kbuf_stride_h = k_cache.data_ptr(), k_cache.stride(0), k_cache.stride(1)
    p.v_buf, p.vbuf_stride_t, p.vbuf_stride_h = v_cache.data_ptr(), v_cache.stride(0), v_cache.stride(1)
    p.out = output.data_ptr()
    p.req_to_token, p.req_to_token_stride = req_to_token.data_ptr(), req_to_token.stride(0)
    p.idx_is_i64 = int(req_to_token.dtype == torch.int64)
    p.req_pool_indices = req_pool_indices.data_ptr() if req_pool_indices is not None else None
    p.seq_lens = seq_lens.data_ptr()
    p.meta_i32_mask = (int(seq_lens.dtype == torch.int32)
                       | (int(req_pool_indices is not None and req_pool_indices.dtype == torch.int32) << 1)
                       | (int(key is not None and loc.dtype == torch.int32) << 2))
    p.page_size = int(page_size)
    if q_rope is not None:
        p.q_rope, p.qrope_stride_b, p.qrope_stride_h = q_rope.data_ptr(), q_rope.stride(0), q_rope.stride(1)
    if key is not None and k_rope is not None:
        p.k_rope_new, p.krope_stride_b = k_rope.data_ptr(), k_rope.stride(0)
    p.B, p.Hq, p.Hkv, p.D, p.Dv = B, Hq, Hkv, D, Dv
    p.dtype = DTYPE_CODE[query.dtype]
    p.write_kv = write_kv
    p.max_seq_len = int(max_seq_len)
    p.sm_scale, p.logit_cap = float(sm_scale), float(logit_cap)
    if kv_fp8:
        p.kv_dtype = 3   # B200_KV_E4M3
        p.k_scale = float(k_scale) if k_scale is not None else 1.0
        p.v_scale = float(v_scale) if v_scale is not None else 1.0
    if cascade is not None:
        cascade.bind(p, B, Hq, Dv, query.dtype)
    S = lib.b200_decode_attn_num_splits(B, Hq, Hkv, D, Dv, int(max_seq_len), int(num_splits)) if B > 0 else 1
    p.num_splits = S
    need = lib.b200_decode_attn_workspace_size(B, Hq, Dv, S)
    if attn_logits is None or attn_logits.numel() * attn_logits.element_size() < need or not attn_logits.is_contiguous():
        attn_logits = torch.empty(need, dtype=torch.uint8, device=dev)
    p.workspace, p.workspace_bytes = attn_logits.data_ptr(), attn_logits.numel() * attn_logits.element_size()
    with torch.cuda.device(dev):
        if not quant_fp8:
            check(lib.b200_decode_attn(p, current_stream(dev)))
            return output
        x_q, x_s = _fp8_group_outputs(B, Hq * Dv, dev)
        rc = _lib.B200_ERR_UNSUPPORTED
        if B > 0 and Dv == 128:
            rc = lib.b200_decode_attn_quant_fp8(p, ptr(x_q), Hq * Dv, ptr(x_s), x_s.stride(0), x_s.stride(1), quant_eps,
                                                -FP8_MAX, FP8_MAX, current_stream(dev))
        if rc == _lib.B200_ERR_UNSUPPORTED:  # shape outside the fused epilogue: same result from two launches
            check(lib.b200_decode_attn(p, current_stream(dev)))
            x_q, x_s = per_token_group_quant_fp8(output.view(B, Hq * Dv), 128, eps=quant_eps, column_major_scales=True)
        else:
            check(rc)
    return output, x_q, x_s


class CascadeState:
    """Device-side state of cascade (shared-prefix) decode for one batch: the shared prefix length P and the one-request
    view of the batch (b200_shared_prefix_len, once per step), scratch for the suffix states, and -- per layer -- the
    prefix state (prefix_o, prefix_lse) written by extend_attention(prefix_only=True).  Everything lives in static
    buffers and P stays on the device, so a captured CUDA graph follows the batch."""

    def __init__(self, max_bs: int, Hq: int, Dv: int, dtype: torch.dtype, device, align: int = 128, min_len: int = 256):
        self.max_bs, self.Hq, self.Dv, self.dtype, self.align, self.min_len = max_bs, Hq, Dv, dtype, align, min_len
        self.scratch = torch.full((1,), 0x7FFFFFFF, dtype=torch.int32, device=device)
        self.prefix_len = torch.zeros(1, dtype=torch.int32, device=device)
        self.one_seq_len = torch.zeros(1, dtype=torch.int64, device=device)
        self.one_req = torch.zeros(1, dtype=torch.int64, device=device)
        self.one_start = torch.zeros(1, dtype=torch.int64, device=device)
        self.ext_lens = {}     # batch size -> int64 [1] device tensor (extend_seq_lens of the one-request view)
        self.casc_o = torch.empty(max_bs, Hq, Dv, dtype=torch.float32, device=device)
        self.casc_lse = torch.empty(max_bs, Hq, dtype=torch.float32, device=device)
        self.prefix_o = torch.empty(max_bs, Hq, Dv, dtype=torch.float32, device=device)
        self.prefix_lse = torch.empty(max_bs, Hq, dtype=torch.float32, device=device)

    def detect(self, req_to_token, req_pool_indices, seq_lens) -> None:
        """Once per decode step (before the layers): P = shared page-table prefix of the batch, on the device."""
        B = seq_lens.numel()
        dev = seq_lens.device
        meta = int(seq_lens.dtype == torch.int32) | (int(req_pool_indices is not None and req_pool_indices.dtype == torch.int32) << 1)
        with torch.cuda.device(dev):
            check(load().b200_shared_prefix_len(ptr(req_to_token), int(req_to_token.dtype == torch.int64), req_to_token.stride(0),
                                                ptr(req_pool_indices), ptr(seq_lens), meta, B, self.align, self.min_len,
                                                ptr(self.scratch), ptr(self.prefix_len), ptr(self.one_seq_len), ptr(self.one_req),
                                                current_stream(dev)))

    def prefix_attention(self, query, k_cache, v_cache, req_to_token, sm_scale: float, logit_cap: float) -> None:
        """Per layer: the state of every (request, head) over the shared prefix, computed ONCE for the batch -- the B
        queries are the `new tokens` of one request whose prefix is the shared one (tuned extend kernel, prefix_only)."""
        B = query.shape[0]
        if B not in self.ext_lens:
            self.ext_lens[B] = torch.full((1,), B, dtype=torch.int64, device=query.device)
        extend_attention(query, query[:, :k_cache.shape[1]], query[:, :k_cache.shape[1]], None, k_cache, v_cache,
                         req_to_token, self.one_req, self.one_seq_len, self.ext_lens[B], self.one_start, B, sm_scale, logit_cap,
                         prefix_only=True, lse_out=self.prefix_lse[:B], out_f32=self.prefix_o[:B])

    def bind(self, p, B: int, Hq: int, Dv: int, dtype) -> None:
        if B > self.max_bs or Hq != self.Hq or Dv != self.Dv or dtype != self.dtype:
            raise RuntimeError("CascadeState was sized for another batch / head shape")
        p.kv_start = self.prefix_len.data_ptr()
        p.cascade_o, p.cascade_lse = self.casc_o.data_ptr(), self.casc_lse.data_ptr()
        p.prefix_o, p.prefix_lse = self.prefix_o.data_ptr(), self.prefix_lse.data_ptr()


def merge_state(v_a: torch.Tensor, s_a: torch.Tensor, v_b: torch.Tensor, s_b: torch.Tensor,
                v_merged: Optional[torch.Tensor] = None, s_merged: Optional[torch.Tensor] = None):
    """sgl_kernel.merge_state (sgl-kernel/python/sgl_kernel/attention.py:12-28; csrc/attention/cascade.cu:12-55): the
    attention states (V [T, H, D], S = natural-log LSE [T, H]) of the same queries over two disjoint key sets -> the
    state over their union.  Returns (v_merged, s_merged)."""
    dev = require_cuda(v_a, s_a, v_b, s_b, v_merged, s_merged)
    if v_a.dim() != 3 or s_a.dim() != 2 or v_a.shape != v_b.shape or s_a.shape != s_b.shape or v_a.shape[:2] != s_a.shape:
        raise RuntimeError("merge_state: v_a / v_b must be [T, H, D] and s_a / s_b [T, H]")   # cascade.cu:22-31
    if v_a.dtype not in DTYPE_CODE or v_a.dtype == torch.float32 or v_b.dtype != v_a.dtype:
        raise RuntimeError("merge_state: v_a / v_b must share a bf16/f16 dtype")
    s_a, s_b = s_a.to(torch.float32).contiguous(), s_b.to(torch.float32).contiguous()
    v_a, v_b = v_a.contiguous(), v_b.contiguous()
    if v_merged is None:
        v_merged = torch.empty_like(v_a)
    if s_merged is None:
        s_merged = torch.empty_like(s_a)
    T, H, D = v_a.shape
    with torch.cuda.device(dev):
        check(load().b200_merge_state(ptr(v_a), ptr(s_a), ptr(v_b), ptr(s_b), ptr(v_merged), ptr(s_merged), T, H, D,
                                      DTYPE_CODE[v_a.dtype], current_stream(dev)))
    return v_merged, s_merged


merge_state_v2 = merge_state   # attention.py:31-52: same contract, another kernel in the reference


def cutlass_mla_get_workspace_size(max_seq_len: int, num_batches: int, sm_count: int = 0, num_kv_splits: int = 1) -> int:
    """sgl_kernel.cutlass_mla_get_workspace_size (sgl-kernel/python/sgl_kernel/attention.py:128-138;
    cutlass_mla_kernel.cu:248-266).  The tuned kernel's scratch does not grow with the batch or the context."""
    assert max_seq_len > 0, f"max_seq_len must be greater than 0, got {max_seq_len}"
    assert num_batches > 0, f"num_batches must be greater than 0, got {num_batches}"
    return int(load().b200_decode_attn_workspace_size(num_batches, 128, 512, 1))


def cutlass_mla_decode(q_nope: torch.Tensor, q_pe: torch.Tensor, kv_c_and_k_pe_cache: torch.Tensor, seq_lens: torch.Tensor,
                       page_table: torch.Tensor, workspace: torch.Tensor, sm_scale: float, num_kv_splits: int = 1) -> torch.Tensor:
    """sgl_kernel.cutlass_mla_decode (sgl-kernel/python/sgl_kernel/attention.py:54-125 ->
    csrc/attention/cutlass_mla_kernel.cu:210-246): MLA absorbed decode over a paged latent cache.

    q_nope [B,H,512], q_pe [B,H,64] (any batch/head strides), kv_c_and_k_pe_cache [pages, P, 576], seq_lens int32 [B],
    page_table int32 [B, blocks] of page ids -> [B, H, 512].  Same asserts as the reference wrapper; no padding of H to
    128 (the kernel works on 16-head blocks) and no concatenation of q_nope / q_pe (two pointers)."""
    assert q_nope.ndim == 3, f"q_nope must be a 3D tensor, but got {q_nope.ndim}"
    assert q_pe.ndim == 3, f"q_pe must be a 3D tensor, but got {q_pe.ndim}"
    assert kv_c_and_k_pe_cache.ndim == 3, f"kv_c_and_k_pe_cache must be a 3D tensor, but got {kv_c_and_k_pe_cache.ndim}"
    B_q, H, D_q_nope = q_nope.shape
    B_q_2, H_2, D_q_pe = q_pe.shape
    assert (B_q == B_q_2) and (H == H_2)
    _, PAGE_SIZE, D_ckv = kv_c_and_k_pe_cache.shape
    assert D_q_nope == 512 and D_q_pe == 64 and D_ckv == 576
    assert H <= 128, f"H must be <= 128, but got {H}"
    assert len(page_table.shape) == 2
    B_block_table, block_num = page_table.shape
    assert B_block_table == B_q
    assert block_num > 0, f"block num must be greater than 0, got {block_num}"
    assert block_num % (128 / PAGE_SIZE) == 0
    assert q_nope.dtype in (torch.float16, torch.bfloat16), f"q_nope.dtype needs to be fp16 or bf16 but got {q_nope.dtype}."
    assert q_nope.dtype == q_pe.dtype == kv_c_and_k_pe_cache.dtype
    assert seq_lens.dtype == torch.int32, f"seq_lens.dtype needs to be int32 but got {seq_lens.dtype}."
    assert page_table.dtype == torch.int32, f"page_table.dtype needs to be int32 but got {page_table.dtype}."
    if PAGE_SIZE & (PAGE_SIZE - 1):
        raise NotImplementedError("cutlass_mla_decode: page size must be a power of two")
    out = q_nope.new_empty((B_q, H, 512))
    if not kv_c_and_k_pe_cache.is_contiguous():
        raise RuntimeError("kv_c_and_k_pe_cache must be contiguous")
    rows = kv_c_and_k_pe_cache.view(-1, 1, 576)
    if page_table.stride(1) != 1:
        page_table = page_table.contiguous()
    decode_attention(q_nope, rows, rows[:, :, :512], out, None, None, None, workspace, page_table, None,
                     seq_lens.contiguous(), sm_scale, 0.0, q_rope=q_pe, page_size=PAGE_SIZE)
    return out


def extend_attention(q_extend, k_extend, v_extend, o_extend, k_buffer, v_buffer, req_to_token, req_pool_indices, seq_lens,
                     extend_seq_lens, extend_start_loc, max_len_extend: int, sm_scale: float, logit_cap: float, *,
                     prefix_only: bool = False, lse_out: Optional[torch.Tensor] = None, out_f32: Optional[torch.Tensor] = None):
    """Same positional signature as torch.ops.sgl_kernel.extend_attention_cpu (torch_extension_cpu.cpp:271-275).
    Keyword-only extras (no reference counterpart): prefix_only -- every new token attends to exactly the prefix
    (the shared-prefix half of cascade decode); lse_out -- float32 [T, Hq] natural-log LSE per row for merge_state;
    out_f32 -- float32 [T, Hq, Dv] output instead of o_extend (which may be None), softmax weights kept exact."""
    dev = require_cuda(q_extend, k_extend, v_extend, o_extend, k_buffer, v_buffer, req_to_token, req_pool_indices,
                       seq_lens, extend_seq_lens, extend_start_loc)
    for t in (q_extend, k_extend, v_extend, k_buffer, v_buffer):
        if t.dim() != 3 or t.stride(-1) != 1:
            raise RuntimeError("q/k/v tensors must be 3-d and contiguous in the last dimension")
    T, Hq, D = q_extend.shape
    Hkv, Dv = k_buffer.shape[1], v_buffer.shape[2]
    if o_extend is None and out_f32 is None:
        raise RuntimeError("o_extend (or out_f32) is required")
    if o_extend is not None and (tuple(o_extend.shape) != (T, Hq, Dv) or not o_extend.is_contiguous()):
        raise RuntimeError("o_extend must be a contiguous [T, Hq, Dv] tensor")
    if out_f32 is not None and (out_f32.dtype != torch.float32 or tuple(out_f32.shape) != (T, Hq, Dv) or not out_f32.is_contiguous()):
        raise RuntimeError("out_f32 must be a contiguous float32 [T, Hq, Dv] tensor")
    rpi, sl, el, es = _i64(req_pool_indices), _i64(seq_lens), _i64(extend_seq_lens), _i64(extend_start_loc)
    p = _lib.ExtendAttnParams()
    p.q, p.q_stride_t, p.q_stride_h = q_extend.data_ptr(), q_extend.stride(0), q_extend.stride(1)
    p.k_ext, p.kext_stride_t, p.kext_stride_h = k_extend.data_ptr(), k_extend.stride(0), k_extend.stride(1)
    p.v_ext, p.vext_stride_t, p.vext_stride_h = v_extend.data_ptr(), v_extend.stride(0), v_extend.stride(1)
    p.out = o_extend.data_ptr() if o_extend is not None else None
    if out_f32 is not None:
        p.out_f32 = out_f32.data_ptr()
    p.k_buf, p.kbuf_stride_t, p.kbuf_stride_h = k_buffer.data_ptr(), k_buffer.stride(0), k_buffer.stride(1)
    p.v_buf, p.vbuf_stride_t, p.vbuf_stride_h = v_buffer.data_ptr(), v_buffer.stride(0), v_buffer.stride(1)
    p.req_to_token, p.req_to_token_stride = req_to_token.data_ptr(), req_to_token.stride(0)
    p.idx_is_i64 = int(req_to_token.dtype == torch.int64)
    p.req_pool_indices, p.seq_lens = rpi.data_ptr(), sl.data_ptr()
    p.extend_seq_lens, p.extend_start_loc = el.data_ptr(), es.data_ptr()
    p.B, p.Hq, p.Hkv, p.D, p.Dv = rpi.numel(), Hq, Hkv, D, Dv
    p.max_len_extend = int(max_len_extend)
    p.dtype = DTYPE_CODE[q_extend.dtype]
    p.sm_scale, p.logit_cap = float(sm_scale), float(logit_cap)
    p.prefix_only = int(bool(prefix_only))
    if lse_out is not None:
        if lse_out.dtype != torch.float32 or tuple(lse_out.shape) != (T, Hq) or not lse_out.is_contiguous():
            raise RuntimeError("lse_out must be a contiguous float32 [T, Hq] tensor")
        p.lse_out = lse_out.data_ptr()
    with torch.cuda.device(dev):
        check(load().b200_extend_attn(p, current_stream(dev)))
    return o_extend


def set_kv_buffer(k_buffer, v_buffer, loc, cache_k, cache_v) -> None:
    """KVCache.set_kv_buffer's scatter (memory_pool.py:454-455): k_buffer[loc] = cache_k; v_buffer[loc] = cache_v."""
    dev = require_cuda(k_buffer, v_buffer, loc, cache_k, cache_v)
    loc = _i64(loc)
    B = loc.numel()
    ck = cache_k.view(B, k_buffer.shape[1], k_buffer.shape[2])
    cv = cache_v.view(B, v_buffer.shape[1], v_buffer.shape[2]) if cache_v is not None else None
    with torch.cuda.device(dev):
        check(load().b200_set_kv_buffer(
            ptr(k_buffer), k_buffer.stride(0), k_buffer.stride(1), ptr(v_buffer), v_buffer.stride(0), v_buffer.stride(1),
            ptr(ck), ck.stride(0), ck.stride(1), ptr(cv), cv.stride(0) if cv is not None else 0,
            cv.stride(1) if cv is not None else 0, ptr(loc), B, k_buffer.shape[1], k_buffer.shape[2], v_buffer.shape[2],
            k_buffer.element_size(), current_stream(dev)))


def set_kv_buffer_fp8(k_buffer, v_buffer, loc, cache_k, cache_v, k_scale: Optional[float] = None,
                      v_scale: Optional[float] = None) -> None:
    """KVCache.set_kv_buffer into an FP8 pool (memory_pool.py:431-455): buffer[loc] = e4m3(cache / scale).
    k_buffer / v_buffer: float8_e4m3fn (or the uint8 storage) [T, Hkv, D]; cache_k / cache_v: bf16 / f16."""
    dev = require_cuda(k_buffer, v_buffer, loc, cache_k, cache_v)
    if k_buffer.element_size() != 1 or cache_k.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("set_kv_buffer_fp8: an 8-bit pool and bf16/f16 rows are required")
    loc = _i64(loc)
    B = loc.numel()
    ck = cache_k.view(B, k_buffer.shape[1], k_buffer.shape[2])
    cv = cache_v.view(B, v_buffer.shape[1], v_buffer.shape[2])
    with torch.cuda.device(dev):
        check(load().b200_set_kv_buffer_fp8(
            ptr(k_buffer), k_buffer.stride(0), k_buffer.stride(1), ptr(v_buffer), v_buffer.stride(0), v_buffer.stride(1),
            ptr(ck), ck.stride(0), ck.stride(1), ptr(cv), cv.stride(0), cv.stride(1), ptr(loc), B, k_buffer.shape[1],
            k_buffer.shape[2], v_buffer.shape[2], DTYPE_CODE[cache_k.dtype], float(k_scale or 1.0), float(v_scale or 1.0),
            current_stream(dev)))


# ---------------------------------------------------------------------------------------------
# the steps either side of the path
# ---------------------------------------------------------------------------------------------
def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    dev = require_cuda(x, weight, residual)
    assert x.is_contiguous() and x.dim() == 2
    out = torch.empty_like(x)
    with torch.cuda.device(dev):
        check(load().b200_rmsnorm(ptr(x), ptr(residual), ptr(weight), ptr(out), DTYPE_CODE[x.dtype], x.shape[0], x.shape[1],
                                  eps, current_stream(dev)))
    return out


def _fp8_group_outputs(M: int, K: int, dev) -> Tuple[torch.Tensor, torch.Tensor]:
    """(x_q [M,K] e4m3, x_s [M,K/128] f32 with stride(0) == 1) -- the operands fp8_blockwise_scaled_mm wants,
    as per_token_group_quant_fp8(column_major_scales=True) lays them out."""
    x_q = torch.empty((M, K), dtype=fp8_dtype, device=dev)
    x_s = torch.empty((K // 128, M), dtype=torch.float32, device=dev).t()
    return x_q, x_s


def rmsnorm_quant_fp8(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None,
                      quant_eps: float = 1e-10, want_half: bool = False):
    """RMSNorm (+ in-place residual add, layernorm.py forward_cuda / fused_add_rmsnorm) fused with
    sglang_per_token_group_quant_fp8(group 128, column-major scales): returns (x_q, x_s[, normed]) bit-identical to
    rmsnorm() followed by per_token_group_quant_fp8()."""
    dev = require_cuda(x, weight, residual)
    assert x.is_contiguous() and x.dim() == 2 and x.shape[1] % 128 == 0
    M, H = x.shape
    x_q, x_s = _fp8_group_outputs(M, H, dev)
    out = torch.empty_like(x) if want_half else None
    if M > 0:
        with torch.cuda.device(dev):
            check(load().b200_rmsnorm_quant_fp8(ptr(x), ptr(residual), ptr(weight), ptr(out), ptr(x_q), H, ptr(x_s),
                                                x_s.stride(0), x_s.stride(1), DTYPE_CODE[x.dtype], M, H, eps, quant_eps,
                                                -FP8_MAX, FP8_MAX, current_stream(dev)))
    return (x_q, x_s, out) if want_half else (x_q, x_s)


def silu_and_mul_quant_fp8(x: torch.Tensor, quant_eps: float = 1e-10):
    """silu_and_mul (activation.py SiluAndMul.forward_cuda) fused with the group-128 FP8 quantiser: (x_q, x_s)
    bit-identical to silu_and_mul() followed by per_token_group_quant_fp8(column_major_scales=True)."""
    dev = require_cuda(x)
    assert x.is_contiguous() and x.dim() == 2 and (x.shape[1] // 2) % 128 == 0
    M, d = x.shape[0], x.shape[1] // 2
    x_q, x_s = _fp8_group_outputs(M, d, dev)
    if M > 0:
        with torch.cuda.device(dev):
            check(load().b200_silu_mul_quant_fp8(ptr(x), ptr(x_q), d, ptr(x_s), x_s.stride(0), x_s.stride(1),
                                                 DTYPE_CODE[x.dtype], M, d, quant_eps, -FP8_MAX, FP8_MAX, current_stream(dev)))
    return x_q, x_s


def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    dev = require_cuda(x)
    assert x.is_contiguous() and x.dim() == 2
    d = x.shape[1] // 2
    out = torch.empty((x.shape[0], d), dtype=x.dtype, device=dev)
    with torch.cuda.device(dev):
        check(load().b200_silu_and_mul(ptr(x), ptr(out), DTYPE_CODE[x.dtype], x.shape[0], d, current_stream(dev)))
    return out


def rope_neox_(q: torch.Tensor, k: torch.Tensor, positions: torch.Tensor, head_dim: int, theta: float = 10000.0) -> None:
    """In-place neox rotary embedding (rotary_embedding.py forward_native, is_neox_style) on q [M, Hq*D] / k [M, Hkv*D];
    q and k may be column slices of a packed qkv tensor (row stride != width)."""
    dev = require_cuda(q, k, positions)
    M = positions.numel()
    q2, k2 = q.view(M, -1) if q.is_contiguous() else q, k.view(M, -1) if k.is_contiguous() else k
    assert q2.dim() == 2 and k2.dim() == 2 and q2.stride(1) == 1 and k2.stride(1) == 1 and positions.dtype == torch.int64
    with torch.cuda.device(dev):
        check(load().b200_rope_neox(ptr(q2), q2.stride(0), ptr(k2), k2.stride(0), ptr(positions), DTYPE_CODE[q.dtype], M,
                                    q2.shape[1] // head_dim, k2.shape[1] // head_dim, head_dim, theta, current_stream(dev)))
