"""ctypes binding of the C-ABI library (include/b200_sgl.h -> csrc/libb200sgl.so).

There is no CPU fallback anywhere in this package: if the library is missing it is built with nvcc
(in-tree), and if that is impossible the import fails loudly; on a device that is not sm_100 every
compute entry point returns B200_ERR_UNSUPPORTED, which surfaces as NotImplementedError.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
SO_PATH = os.environ.get("B200_SGL_LIB") or os.path.join(_CSRC, "libb200sgl.so")   # (override: developer probe builds)

B200_OK, B200_ERR_ARG, B200_ERR_UNSUPPORTED, B200_ERR_CUDA = 0, -1, -2, -3
BF16, F16, F32 = 0, 1, 2
DTYPE_CODE = {torch.bfloat16: BF16, torch.float16: F16, torch.float32: F32}

c_void_p, c_int, c_int64, c_float, c_size_t = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t


class DecodeAttnParams(C.Structure):
    """struct b200_decode_attn_params (include/b200_sgl.h)."""

    _fields_ = [
        ("q", c_void_p), ("q_stride_b", c_int64), ("q_stride_h", c_int64),
        ("k_new", c_void_p), ("knew_stride_b", c_int64), ("knew_stride_h", c_int64),
        ("v_new", c_void_p), ("vnew_stride_b", c_int64), ("vnew_stride_h", c_int64),
        ("k_buf", c_void_p), ("kbuf_stride_t", c_int64), ("kbuf_stride_h", c_int64),
        ("v_buf", c_void_p), ("vbuf_stride_t", c_int64), ("vbuf_stride_h", c_int64),
        ("out", c_void_p),
        ("req_to_token", c_void_p), ("req_to_token_stride", c_int64),
        ("req_pool_indices", c_void_p), ("seq_lens", c_void_p), ("out_cache_loc", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_size_t),
        ("idx_is_i64", C.c_int32),
        ("B", C.c_int32), ("Hq", C.c_int32), ("Hkv", C.c_int32), ("D", C.c_int32), ("Dv", C.c_int32),
        ("dtype", C.c_int32), ("write_kv", C.c_int32), ("num_splits", C.c_int32), ("max_seq_len", C.c_int32),
        ("sm_scale", c_float), ("logit_cap", c_float),
        ("meta_i32_mask", C.c_int32), ("page_size", C.c_int32),
        ("q_rope", c_void_p), ("qrope_stride_b", c_int64), ("qrope_stride_h", c_int64),
        ("k_rope_new", c_void_p), ("krope_stride_b", c_int64),
        ("kv_dtype", C.c_int32), ("k_scale", c_float), ("v_scale", c_float), ("reserved1", C.c_int32),
        ("kv_start", c_void_p), ("cascade_o", c_void_p), ("cascade_lse", c_void_p), ("prefix_o", c_void_p), ("prefix_lse", c_void_p),
    ]


class ExtendAttnParams(C.Structure):
    """struct b200_extend_attn_params (include/b200_sgl.h)."""

    _fields_ = [
        ("q", c_void_p), ("q_stride_t", c_int64), ("q_stride_h", c_int64),
        ("k_ext", c_void_p), ("kext_stride_t", c_int64), ("kext_stride_h", c_int64),
        ("v_ext", c_void_p), ("vext_stride_t", c_int64), ("vext_stride_h", c_int64),
        ("out", c_void_p),
        ("k_buf", c_void_p), ("kbuf_stride_t", c_int64), ("kbuf_stride_h", c_int64),
        ("v_buf", c_void_p), ("vbuf_stride_t", c_int64), ("vbuf_stride_h", c_int64),
        ("req_to_token", c_void_p), ("req_to_token_stride", c_int64),
        ("req_pool_indices", c_void_p), ("seq_lens", c_void_p), ("extend_seq_lens", c_void_p),
        ("extend_start_loc", c_void_p),
        ("idx_is_i64", C.c_int32),
        ("B", C.c_int32), ("Hq", C.c_int32), ("Hkv", C.c_int32), ("D", C.c_int32), ("Dv", C.c_int32),
        ("max_len_extend", C.c_int32), ("dtype", C.c_int32),
        ("sm_scale", c_float), ("logit_cap", c_float),
        ("prefix_only", C.c_int32), ("reserved0", C.c_int32), ("lse_out", c_void_p), ("out_f32", c_void_p),
    ]


# name -> (restype, argtypes); must list EVERY symbol include/b200_sgl.h declares
SIGNATURES = {
    "b200_last_error": (C.c_char_p, []),
    "b200_abi_version": (c_int, []),
    "b200_sm_arch": (c_int, []),
    "b200_sm_count": (c_int, []),
    "b200_launch_count": (C.c_ulonglong, []),
    "b200_tune_set": (c_int, [C.c_char_p, c_int]),
    "b200_tune_get": (c_int, [C.c_char_p, C.POINTER(c_int)]),
    "b200_decode_attn_num_splits": (c_int, [c_int] * 7),
    "b200_decode_attn_workspace_size": (c_size_t, [c_int] * 4),
    "b200_decode_attn": (c_int, [C.POINTER(DecodeAttnParams), c_void_p]),
    "b200_decode_attn_static": (c_int, [C.POINTER(DecodeAttnParams), c_void_p]),
    "b200_decode_attn_quant_fp8": (c_int, [C.POINTER(DecodeAttnParams), c_void_p, c_int64, c_void_p, c_int64, c_int64,
                                           c_float, c_float, c_float, c_void_p]),
    "b200_set_kv_buffer": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64,
                                   c_void_p, c_int64, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_set_kv_buffer_fp8": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64,
                                       c_void_p, c_int64, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float, c_float,
                                       c_void_p]),
    "b200_extend_attn": (c_int, [C.POINTER(ExtendAttnParams), c_void_p]),
    "b200_merge_state": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    "b200_shared_prefix_len": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                       c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200_create_kv_indices": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                       c_void_p]),
    "b200_create_page_table": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                       c_int, c_int, c_void_p]),
    "b200_per_token_group_quant_fp8": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int64, c_int64,
                                               c_int, c_int, c_int, c_float, c_float, c_float, c_void_p]),
    "b200_per_token_quant_int8": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int, c_int, c_void_p]),
    "b200_gemm_workspace_size": (c_size_t, [c_int] * 3),
    "b200_fp8_blockwise_scaled_mm": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_void_p,
                                             c_int64, c_int64, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p,
                                             c_size_t, c_void_p]),
    "b200_fp8_blockwise_scaled_mm_rope": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_void_p,
                                                  c_int64, c_int64, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p,
                                                  c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "b200_rope_cos_sin": (c_int, [c_void_p, c_int, c_int, c_float, c_void_p, c_void_p]),
    "b200_int8_scaled_mm": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                    c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "b200_awq_dequantize": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_awq_w4a16_workspace_size": (c_size_t, [c_int] * 3),
    "b200_awq_w4a16_mm": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                  c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "b200_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p]),
    "b200_rmsnorm_quant_fp8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int64,
                                       c_int, c_int, c_int, c_float, c_float, c_float, c_float, c_void_p]),
    "b200_silu_mul_quant_fp8": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_int, c_int,
                                        c_float, c_float, c_float, c_void_p]),
    "b200_silu_and_mul": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "b200_rope_neox": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float,
                             c_void_p]),
    "b200_ar_region_bytes": (c_size_t, [c_size_t]),
    "b200_ar_alloc": (c_int, [c_size_t, C.POINTER(c_void_p), c_void_p]),
    "b200_ar_init": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t, C.POINTER(c_void_p)]),
    "b200_ar_allreduce": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b200_ar_mode": (c_int, [c_int, c_size_t]),
    "b200_ar_mode2": (c_int, [c_int, c_size_t, c_int]),
    "b200_ar_set_mode": (c_int, [c_void_p, c_int]),
    "b200_ar_free_region": (None, [c_void_p]),
    "b200_ar_init_ptrs": (c_int, [C.POINTER(C.c_uint64), C.c_uint64, c_int, c_int, c_size_t, C.POINTER(c_void_p)]),
    "b200_ar_has_multicast": (c_int, [c_void_p]),
    "b200_ar_allreduce_add_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p,
                                              c_int64, c_int64, c_int, c_int, c_int, c_float, c_float, c_float, c_float,
                                              c_void_p]),
    "b200_ar_error": (c_int, [c_void_p]),
    "b200_ar_destroy": (None, [c_void_p]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load (building first if the sources are newer) the C-ABI library; raises if that is impossible."""
    global _lib
    if _lib is not None:
        return _lib
    from .csrc import build as _build

    if not os.path.exists(SO_PATH) or (os.path.exists(_build.NVCC) and not _build.up_to_date()):
        _build.build()
    if not os.path.exists(SO_PATH):
        raise ImportError(f"{SO_PATH} is missing and could not be built; iaas_sglang_b200 has no CPU fallback")
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.b200_abi_version() != 4:
        raise ImportError("libb200sgl.so ABI version mismatch")
    _lib = lib
    return lib


def tune_set(name: str, value: int) -> None:
    """Developer / tuning knob (csrc/runtime.cu kTune); the environment (B200_*) is read once at load."""
    check(load().b200_tune_set(name.encode(), int(value)))


def tune_get(name: str) -> int:
    v = c_int(0)
    check(load().b200_tune_get(name.encode(), C.byref(v)))
    return v.value


def last_error() -> str:
    return load().b200_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    """Map C-ABI return codes onto the reference's error behaviour (TORCH_CHECK -> RuntimeError,
    TORCH_CHECK_NOT_IMPLEMENTED -> NotImplementedError; fp8_blockwise_gemm_kernel.cu:351-421)."""
    if rc == B200_OK:
        return
    msg = last_error()
    if rc == B200_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)


def ptr(t: Optional[torch.Tensor]) -> c_void_p:
    return c_void_p(0 if t is None else t.data_ptr())


def current_stream(device=None) -> c_void_p:
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(*tensors: torch.Tensor) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("iaas_sglang_b200: expected CUDA tensors (this package has no CPU path)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError("iaas_sglang_b200: tensors are on different devices")
    return dev
