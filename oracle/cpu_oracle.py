"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper over oracle/liboracle.so (plain-C restatement).

All functions take CPU torch tensors (bf16 / fp16 / fp8 / int8 / int32 / int64) and mirror the
reference signatures they restate; see oracle.c for the reference file:line of each.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None

BF16, F16, F32 = 0, 1, 2
_DT = {torch.bfloat16: BF16, torch.float16: F16, torch.float32: F32}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_decode_attention.restype = C.c_int
        _lib.orc_extend_attention.restype = C.c_int
    return _lib


def _p(t: Optional[torch.Tensor]):
    if t is None:
        return C.c_void_p(0)
    assert t.device.type == "cpu"
    return C.c_void_p(t.data_ptr())


def _i64(v):
    return C.c_int64(int(v))


def _as_i64(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.int64).contiguous()


def create_kv_indices(req_to_token, req_pool_indices, lens, kv_indptr, kv_start=None) -> torch.Tensor:
    """utils.py:5-41 create_flashinfer_kv_indices_triton."""
    assert req_to_token.dtype in (torch.int32, torch.int64) and req_to_token.stride(1) == 1
    rpi, ln, ip = _as_i64(req_pool_indices), _as_i64(lens), _as_i64(kv_indptr)
    ks = _as_i64(kv_start) if kv_start is not None else None
    out = torch.zeros(int(ip[-1]), dtype=torch.int32)
    lib().orc_create_kv_indices(_p(req_to_token), C.c_int(req_to_token.dtype == torch.int64),
                                _i64(req_to_token.stride(0)), _p(rpi), _p(ln), _p(ip), _p(ks), _p(out),
                                _i64(rpi.numel()))
    return out


def create_page_table(req_to_token, req_pool_indices, lens, page_size, max_pages, kv_start=None, fill=-1):
    """utils.py:44-94 create_flashmla_kv_indices_triton (+ the -1 fill of cutlass_mla_backend.py:91-96)."""
    rpi, ln = _as_i64(req_pool_indices), _as_i64(lens)
    ks = _as_i64(kv_start) if kv_start is not None else None
    out = torch.full((rpi.numel(), max_pages), fill, dtype=torch.int32)
    lib().orc_create_page_table(_p(req_to_token), C.c_int(req_to_token.dtype == torch.int64),
                                _i64(req_to_token.stride(0)), _p(rpi), _p(ln), _p(ks), _p(out),
                                _i64(out.stride(0)), _i64(rpi.numel()), _i64(page_size))
    return out


def decode_attention(query, k_cache, v_cache, output, key, value, loc, attn_logits, req_to_token,
                     req_pool_indices, seq_lens, sm_scale, logit_cap, *, num_kv_splits: int = 8,
                     write_kv: bool = True, quirk_empty_split: bool = False, is_mla: Optional[bool] = None):
    """Same positional signature as torch.ops.sgl_kernel.decode_attention_cpu
    (sgl-kernel/csrc/cpu/decode.cpp:1375; schema torch_extension_cpu.cpp:264-268).
    attn_logits may be None (scratch is then private)."""
    B, Hq, D = query.shape
    Hkv, Dv = k_cache.shape[1], v_cache.shape[-1]
    assert query.stride(-1) == 1 and k_cache.stride(-1) == 1 and v_cache.stride(-1) == 1
    assert output.is_contiguous() and output.shape == (B, Hq, Dv)
    if is_mla is None:  # decode.cpp:1455
        is_mla = k_cache.data_ptr() == v_cache.data_ptr() and Hkv == 1 and D == Dv + 64
    if attn_logits is not None:
        assert attn_logits.dtype == torch.float32 and attn_logits.is_contiguous()
        num_kv_splits = attn_logits.shape[2]
    if key is None:
        write_kv = False
    loc64 = _as_i64(loc) if loc is not None else None
    assert loc64 is not None or not write_kv
    rpi, sl = _as_i64(req_pool_indices), _as_i64(seq_lens)
    rc = lib().orc_decode_attention(
        _p(query), _i64(query.stride(0)), _i64(query.stride(1)),
        _p(k_cache), _i64(k_cache.stride(0)), _i64(k_cache.stride(1)),
        _p(v_cache), _i64(v_cache.stride(0)), _i64(v_cache.stride(1)),
        _p(output),
        _p(key), _i64(key.stride(0) if key is not None else 0), _i64(key.stride(1) if key is not None else 0),
        _p(value), _i64(value.stride(0) if value is not None else 0), _i64(value.stride(1) if value is not None else 0),
        _p(loc64), _p(attn_logits), _p(req_to_token), C.c_int(req_to_token.dtype == torch.int64),
        _i64(req_to_token.stride(0)), _p(rpi), _p(sl), C.c_int(B), C.c_int(Hq), C.c_int(Hkv), C.c_int(D), C.c_int(Dv),
        C.c_int(num_kv_splits), C.c_float(sm_scale), C.c_float(logit_cap), C.c_int(_DT[query.dtype]),
        C.c_int(bool(is_mla)), C.c_int(bool(write_kv)), C.c_int(bool(quirk_empty_split)))
    if rc != 0:
        raise RuntimeError(f"orc_decode_attention failed rc={rc}")
    return output


def extend_attention(q_extend, k_extend, v_extend, o_extend, k_buffer, v_buffer, req_to_token, req_pool_indices,
                     seq_lens, extend_seq_lens, extend_start_loc, max_len_extend, sm_scale, logit_cap):
    """Same positional signature as torch.ops.sgl_kernel.extend_attention_cpu (torch_extension_cpu.cpp:271-275)."""
    T, Hq, D = q_extend.shape
    Hkv, Dv = k_buffer.shape[1], v_buffer.shape[-1]
    assert o_extend.is_contiguous()
    rpi, sl, el, es = map(_as_i64, (req_pool_indices, seq_lens, extend_seq_lens, extend_start_loc))
    rc = lib().orc_extend_attention(
        _p(q_extend), _i64(q_extend.stride(0)), _i64(q_extend.stride(1)),
        _p(k_extend), _i64(k_extend.stride(0)), _i64(k_extend.stride(1)),
        _p(v_extend), _i64(v_extend.stride(0)), _i64(v_extend.stride(1)), _p(o_extend),
        _p(k_buffer), _i64(k_buffer.stride(0)), _i64(k_buffer.stride(1)),
        _p(v_buffer), _i64(v_buffer.stride(0)), _i64(v_buffer.stride(1)),
        _p(req_to_token), C.c_int(req_to_token.dtype == torch.int64), _i64(req_to_token.stride(0)),
        _p(rpi), _p(sl), _p(el), _p(es), C.c_int(rpi.numel()), C.c_int(Hq), C.c_int(Hkv), C.c_int(D), C.c_int(Dv),
        C.c_float(sm_scale), C.c_float(logit_cap), C.c_int(_DT[q_extend.dtype]))
    if rc != 0:
        raise RuntimeError(f"orc_extend_attention failed rc={rc}")
    return o_extend


def per_token_group_quant_fp8(x: torch.Tensor, group_size: int = 128, eps: float = 1e-10,
                              column_major_scales: bool = True):
    """fp8_kernel.py:273-326 sglang_per_token_group_quant_fp8 -> (q e4m3 [M,K], s f32 [M,K/g])."""
    assert x.dim() == 2 and x.stride(1) == 1
    M, K = x.shape
    q = torch.empty((M, K), dtype=torch.float8_e4m3fn)
    if column_major_scales:
        s = torch.empty((K // group_size, M), dtype=torch.float32).t()
    else:
        s = torch.empty((M, K // group_size), dtype=torch.float32)
    lib().orc_per_token_group_quant_fp8(_p(x), _i64(x.stride(0)), C.c_int(_DT[x.dtype]), _p(q), _p(s),
                                        _i64(s.stride(0)), _i64(s.stride(1)), _i64(M), _i64(K), C.c_int(group_size),
                                        C.c_float(eps), C.c_float(-448.0), C.c_float(448.0))
    return q, s


def per_token_quant_int8(x: torch.Tensor):
    """int8_kernel.py:51-81 per_token_quant_int8 -> (q int8 [M,K], scale f32 [M,1])."""
    assert x.dim() == 2 and x.stride(1) == 1
    M, K = x.shape
    q = torch.empty((M, K), dtype=torch.int8)
    s = torch.empty((M, 1), dtype=torch.float32)
    lib().orc_per_token_quant_int8(_p(x), _i64(x.stride(0)), C.c_int(_DT[x.dtype]), _p(q), _p(s), _i64(M), _i64(K))
    return q, s


def fp8_blockwise_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype):
    """sgl_kernel.fp8_blockwise_scaled_mm (gemm.py:24-31): mat_b is [K,N] column-major."""
    M, K = mat_a.shape
    N = mat_b.shape[1]
    assert mat_a.stride(1) == 1 and mat_b.stride(0) == 1
    out = torch.empty((M, N), dtype=out_dtype)
    lib().orc_fp8_blockwise_mm(_p(mat_a), _i64(mat_a.stride(0)), _p(mat_b), _i64(mat_b.stride(1)),
                               _p(scales_a), _i64(scales_a.stride(0)), _i64(scales_a.stride(1)),
                               _p(scales_b), _i64(scales_b.stride(0)), _i64(scales_b.stride(1)),
                               _p(out), C.c_int(_DT[out_dtype]), _i64(M), _i64(N), _i64(K))
    return out


def int8_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype, bias=None):
    """sgl_kernel.int8_scaled_mm (gemm.py:13-21): mat_b is [K,N] column-major."""
    M, K = mat_a.shape
    N = mat_b.shape[1]
    assert mat_a.stride(1) == 1 and mat_b.stride(0) == 1
    out = torch.empty((M, N), dtype=out_dtype)
    sa = scales_a.reshape(-1).contiguous().float()
    sb = scales_b.reshape(-1).contiguous().float()
    if bias is not None:
        assert bias.dtype == out_dtype
        bias = bias.contiguous()
    lib().orc_int8_scaled_mm(_p(mat_a), _i64(mat_a.stride(0)), _p(mat_b), _i64(mat_b.stride(1)), _p(sa), _p(sb),
                             _p(bias), _p(out), C.c_int(_DT[out_dtype]), _i64(M), _i64(N), _i64(K))
    return out


def awq_dequantize(qweight, scales, qzeros):
    """sgl_kernel.awq_dequantize (gemm.py:7-10)."""
    rows, cols = qweight.shape
    group = rows // scales.shape[0]
    out = torch.empty((rows, cols * 8), dtype=scales.dtype)
    lib().orc_awq_dequantize(_p(qweight.contiguous()), _p(scales.contiguous()), _p(qzeros.contiguous()), _p(out),
                             C.c_int(_DT[scales.dtype]), _i64(rows), _i64(cols), _i64(group))
    return out


def awq_gemm(x, qweight, scales, qzeros, bias=None):
    """AWQLinearMethod.apply (awq.py:186-204): x @ dequant(W) + bias."""
    M, K = x.shape
    N = qweight.shape[1] * 8
    group = K // scales.shape[0]
    out = torch.empty((M, N), dtype=x.dtype)
    lib().orc_awq_gemm(_p(x), _i64(x.stride(0)), _p(qweight.contiguous()), _p(scales.contiguous()),
                       _p(qzeros.contiguous()), _p(bias), _p(out), C.c_int(_DT[x.dtype]), _i64(M), _i64(N), _i64(K),
                       _i64(group))
    return out


def fp8_w8a16_mm(x, weight, weight_scale, block_size, bias, out_dtype):
    """fp8_scaled_mm_cpu semantics (gemm_fp8.cpp:468-548): W [N,K] e4m3, scale [N/bn, K/bk]."""
    M, K = x.shape
    N = weight.shape[0]
    out = torch.empty((M, N), dtype=out_dtype)
    lib().orc_fp8_w8a16_mm(_p(x), _i64(x.stride(0)), _p(weight), _i64(weight.stride(0)), _p(weight_scale),
                           _i64(weight_scale.stride(0)), _i64(weight_scale.stride(1)), C.c_int(block_size[0]),
                           C.c_int(block_size[1]), _p(bias), _p(out), C.c_int(_DT[out_dtype]), _i64(M), _i64(N), _i64(K))
    return out


def rmsnorm(x, w, eps):
    out = torch.empty_like(x)
    lib().orc_rmsnorm(_p(x.contiguous()), _p(w), _p(out), C.c_int(_DT[x.dtype]), _i64(x.shape[0]), _i64(x.shape[1]),
                      C.c_float(eps))
    return out


def silu_and_mul(x):
    M, d2 = x.shape
    out = torch.empty((M, d2 // 2), dtype=x.dtype)
    lib().orc_silu_and_mul(_p(x.contiguous()), _p(out), C.c_int(_DT[x.dtype]), _i64(M), _i64(d2 // 2))
    return out


def f32_to_e4m3(x: torch.Tensor) -> torch.Tensor:
    x = x.contiguous().float()
    y = torch.empty(x.shape, dtype=torch.uint8)
    lib().orc_cvt_f32_to_e4m3(_p(x), _p(y), _i64(x.numel()))
    return y.view(torch.float8_e4m3fn)


def e4m3_to_f32(x: torch.Tensor) -> torch.Tensor:
    x = x.contiguous()
    y = torch.empty(x.shape, dtype=torch.float32)
    lib().orc_cvt_e4m3_to_f32(_p(x), _p(y), _i64(x.numel()))
    return y


def merge_state(v_a, s_a, v_b, s_b):
    """merge_attn_states_kernel (sgl-kernel/csrc/attention/merge_attn_states.cu:30-120; same contract as merge_state,
    cascade.cu:12-55; torch form in sgl-kernel/tests/test_merge_state_v2.py): states over two disjoint key sets ->
    the state over their union.  v [T, H, D], s = natural-log LSE [T, H]; an infinite s marks an empty state."""
    la, lb = s_a.float().clone(), s_b.float().clone()
    la[torch.isinf(la)] = float("-inf")
    lb[torch.isinf(lb)] = float("-inf")
    mx = torch.maximum(la, lb)
    ea, eb = torch.exp(la - mx), torch.exp(lb - mx)
    den = ea + eb
    wa, wb = ea / den, eb / den
    dead = torch.isinf(mx)
    wa[dead], wb[dead] = 0.0, 0.0
    out = (v_a.float() * wa[..., None] + v_b.float() * wb[..., None]).to(v_a.dtype)
    lse = torch.log(den) + mx
    lse[dead] = float("-inf")
    return out, lse


def shared_prefix_len(req_to_token, req_pool_indices, seq_lens, align: int = 128, min_len: int = 256) -> int:
    """Restatement of b200_shared_prefix_len (no reference counterpart: the reference knows the shared prefix from its
    radix tree): leading page-table entries common to all requests, at most min(seq_len) - 1, rounded down to `align`."""
    B = seq_lens.numel()
    if B < 2:
        return 0
    rows = req_to_token[req_pool_indices.long()]
    lim = int(seq_lens.min()) - 1
    if lim <= 0:
        return 0
    same = (rows[:, :lim] == rows[0:1, :lim]).all(dim=0)
    p = int(same.long().cumprod(0).sum())
    p = p // align * align
    return p if p >= min_len else 0
