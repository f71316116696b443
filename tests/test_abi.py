"""CPU tier: the C-ABI library builds, loads and exports every symbol include/b200_sgl.h declares; error
behaviour of the host layer without a GPU (no CPU fallback: CPU tensors are rejected, compute on a non-sm_100
device is NotImplementedError)."""
import ctypes
import os
import re

import pytest
import torch

from iaas_sglang_b200 import _lib, ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200_sgl.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    syms = declared_symbols()
    assert len(syms) >= 20
    assert set(syms) == set(_lib.SIGNATURES), set(syms) ^ set(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    raw = ctypes.CDLL(_lib.SO_PATH)
    for s in declared_symbols():
        assert hasattr(raw, s), s
    assert lib.b200_abi_version() == 4


def test_struct_layout_matches_header():
    # v1: 23 pointer/int64/size_t fields (8 B) + 10 int32 + 2 float = 232 bytes; v2 adds 2 int32 + 5 pointer/int64
    assert ctypes.sizeof(_lib.DecodeAttnParams) == 23 * 8 + 12 * 4 + 2 * 4 + 5 * 8 + 4 * 4 + 5 * 8   # v3: kv_dtype / scales, five cascade pointers
    assert ctypes.sizeof(_lib.ExtendAttnParams) == 22 * 8 + 10 * 4 + 2 * 4 + 16   # v3: prefix_only, reserved, lse_out, out_f32


def test_size_queries_need_no_gpu():
    lib = _lib.load()
    s = lib.b200_decode_attn_num_splits(64, 32, 8, 128, 128, 4096, 0)
    assert 1 <= s <= 64
    assert lib.b200_decode_attn_workspace_size(64, 32, 128, s) >= 64 * 32 * s * 129 * 4
    # the tuned GQA path partitions KV on the device (always 1); the hint applies to the generic path
    assert lib.b200_decode_attn_num_splits(64, 32, 8, 128, 128, 4096, 5) == 1
    assert lib.b200_decode_attn_num_splits(64, 32, 8, 96, 96, 4096, 5) == 5
    assert lib.b200_gemm_workspace_size(64, 4096, 4096) > 16
    assert lib.b200_awq_w4a16_workspace_size(16, 1024, 512) >= 256  # fused decode path: no weight staging
    assert lib.b200_awq_w4a16_workspace_size(1024, 1024, 512) >= 1024 * 512 * 2  # two-phase path (M > 256)


def test_cpu_tensors_are_rejected():
    a = torch.zeros(4, 128, dtype=torch.float8_e4m3fn)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.fp8_blockwise_scaled_mm(a, a.t(), torch.ones(4, 1), torch.ones(1, 1), torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.awq_dequantize(torch.zeros(128, 16, dtype=torch.int32), torch.ones(1, 128, dtype=torch.float16),
                           torch.zeros(1, 16, dtype=torch.int32))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU error path")
def test_compute_entry_points_fail_loudly_without_sm100():
    lib = _lib.load()
    p = _lib.DecodeAttnParams()
    dummy = ctypes.create_string_buffer(64)
    addr = ctypes.addressof(dummy)
    for f in ("q", "k_buf", "v_buf", "out", "req_to_token", "req_pool_indices", "seq_lens"):
        setattr(p, f, addr)
    p.B, p.Hq, p.Hkv, p.D, p.Dv = 1, 8, 8, 64, 64
    rc = lib.b200_decode_attn(p, None)
    assert rc == _lib.B200_ERR_UNSUPPORTED
    assert b"no fallback" in lib.b200_last_error()
    with pytest.raises(NotImplementedError):
        _lib.check(rc)
    rc = lib.b200_fp8_blockwise_scaled_mm(addr, 128, addr, 128, addr, 1, 1, addr, 1, 1, addr, 128, 0, 1, 128, 128, None, 0, None)
    assert rc == _lib.B200_ERR_UNSUPPORTED
    # argument errors are reported before the device check, like the reference's TORCH_CHECKs
    rc = lib.b200_int8_scaled_mm(addr, 100, addr, 100, addr, addr, None, addr, 128, 0, 1, 128, 100, None, 0, None)
    assert rc == _lib.B200_ERR_ARG and b"multiple of 16" in lib.b200_last_error()
