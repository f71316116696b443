"""Generate the golden vectors under tests/golden/ from the REFERENCE ITSELF.  Run in the authoring
container only (needs /root/reference and oracle/_ref); the fixtures it writes are committed and are what
travels (the GPU box has no /root/reference).

    python tests/golden/make_golden.py

Two sources, both the reference's own code:
  (1) oracle/_ref/libsgl_ref_cpu.so -- the reference's CPU kernels compiled unmodified from
      /root/reference/sgl-kernel/csrc/cpu (oracle/build_ref.py): decode_attention_cpu (GQA/MHA/MLA),
      extend_attention_cpu, fp8_scaled_mm_cpu, int8_scaled_mm_with_quant, per_token_quant_int8_cpu;
  (2) the reference's Python test helpers imported from their files (with `sgl_kernel`/`sglang` stubbed,
      they are plain torch): baseline_scaled_mm (sgl-kernel/tests/test_fp8_blockwise_gemm.py:19-59),
      torch_scaled_mm (tests/test_int8_gemm.py:10-16), awq_dequantize_torch (tests/test_awq_dequant.py:28-57),
      native_per_token_group_quant_fp8 / native_w8a8_block_fp8_matmul (python/sglang/test/test_block_fp8.py:23-48,
      273-325), per_token_quant_int8 / convert_weight / native_w8a8_per_token_matmul (test/srt/cpu/utils.py),
      and the kv-indices reference expression of test/srt/test_create_kvindices.py:43-50.
Inputs follow the reference tests' distributions and seeds (torch.manual_seed(1234)); shapes are the
reference tests' shapes with the context shortened so the fixtures stay small.
"""
from __future__ import annotations

import importlib.util
import os
import sys
from unittest import mock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("B200_SGL_REFERENCE", "/root/reference")


def load_ref_module(relpath: str, name: str):
    """Import a reference test file as a module with sgl_kernel / sglang / triton-decorated bits stubbed."""
    stubs = {}
    for m in ("sgl_kernel", "sglang", "sglang.srt", "sglang.srt.utils", "sglang.test", "sglang.test.test_utils",
              "sglang.srt.layers", "sglang.srt.layers.activation", "sglang.srt.layers.moe",
              "sglang.srt.layers.moe.fused_moe_triton", "sglang.srt.layers.moe.fused_moe_triton.fused_moe",
              "sglang.srt.layers.quantization", "sglang.srt.layers.quantization.fp8_kernel",
              "sglang.srt.layers.quantization.fp8_utils", "sglang.srt.layers.attention",
              "sglang.srt.layers.attention.utils"):
        if m not in sys.modules:
            stubs[m] = mock.MagicMock(name=m)
    tu = stubs.get("sglang.test.test_utils")
    if tu is not None:
        import unittest
        tu.CustomTestCase = unittest.TestCase
    with mock.patch.dict(sys.modules, stubs):
        sys.path.insert(0, os.path.join(REF, "test", "srt", "cpu"))
        try:
            spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
        finally:
            sys.path.pop(0)
    return mod


def u16(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


def u8(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.uint8).numpy()


def save(name: str, **arrays):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(f"{name}.npz  {os.path.getsize(path) / 1024:.0f} KiB")


# ------------------------------------------------------------------------------------------------
def gen_decode(ref):
    """test/srt/cpu/test_decode.py:62-143 (11 shapes) and test_mla.py:67-150, shortened contexts."""
    torch.manual_seed(1234)
    cases = [  # (B, Hq, Hkv, D, Dv, seq_len, mla, logit_cap)
        (2, 16, 16, 64, 64, 40, False, 0.0), (2, 16, 1, 16, 16, 100, False, 0.0), (2, 32, 8, 33, 55, 37, False, 0.0),
        (2, 16, 1, 64, 64, 129, False, 0.0), (2, 64, 1, 13, 13, 64, False, 0.0), (2, 128, 1, 80, 80, 50, False, 0.0),
        (2, 128, 2, 512, 512, 24, False, 0.0), (1, 16, 1, 576, 512, 72, True, 0.0), (1, 16, 16, 576, 512, 9, False, 0.0),
        (1, 22, 1, 576, 512, 64, True, 0.0), (1, 40, 8, 128, 128, 100, False, 0.0), (3, 22, 1, 576, 512, 40, True, 0.0),
        (2, 32, 8, 128, 128, 77, False, 30.0), (4, 32, 8, 128, 128, 5, False, 0.0),  # seq_len < 8 splits
    ]
    out = {}
    for i, (B, Hq, Hkv, D, Dv, S, mla, cap) in enumerate(cases):
        T = B * S
        q = torch.randn(B, Hq, D, dtype=torch.bfloat16)
        k_buf = torch.randn(T, Hkv, D, dtype=torch.bfloat16)
        v_buf = k_buf.narrow(2, 0, Dv) if mla else torch.randn(T, Hkv, Dv, dtype=torch.bfloat16)
        key = torch.randn(B, Hkv, D, dtype=torch.bfloat16)
        value = key.narrow(2, 0, Dv) if mla else torch.randn(B, Hkv, Dv, dtype=torch.bfloat16)
        perm = torch.randperm(T)
        r2t = perm.view(B, S).to(torch.int32)       # shuffled page table (page_size 1)
        seq_lens = torch.randint(max(1, S // 2), S + 1, (B,), dtype=torch.int64)
        seq_lens[0] = S
        rpi = torch.arange(B, dtype=torch.int64)
        loc = torch.stack([r2t[b, seq_lens[b] - 1] for b in range(B)]).to(torch.int64)
        kb2 = k_buf.clone()
        vb2 = kb2.narrow(2, 0, Dv) if mla else v_buf.clone()
        o = torch.zeros(B, Hq, Dv, dtype=torch.bfloat16)
        # zero-initialised like the reference backend allocates it (intel_amx_backend.py:36-46): an EMPTY kv split
        # (seq_len < ~57) then enters the reference's merge with lse = 0 -- see oracle.c quirk_empty_split
        logits = torch.zeros(B, Hq, 8, Dv + 1, dtype=torch.float32)
        ref.decode_attention_cpu(q, kb2, vb2, o, key, value, loc, logits, r2t, rpi, seq_lens, 1.0 / D ** 0.5, cap)
        p = f"c{i}_"
        out.update({p + "shape": np.array([B, Hq, Hkv, D, Dv, S, int(mla)]), p + "cap": np.float32(cap), p + "q": u16(q),
                    p + "k_buf": u16(k_buf), p + "key": u16(key), p + "r2t": r2t.numpy(), p + "seq_lens": seq_lens.numpy(),
                    p + "loc": loc.numpy(), p + "o": u16(o)})
        if not mla:
            out.update({p + "v_buf": u16(v_buf), p + "value": u16(value)})
    out["n_cases"] = np.array(len(cases))
    save("decode_attention_cpu", **out)


def gen_extend(ref):
    """test/srt/cpu/test_extend.py:77-190 (shapes (1,123,1,1,128,96), (1,123,16,1,..), (4,1230->123,16,4,..))."""
    torch.manual_seed(1234)
    out = {}
    cases = [(1, 60, 1, 1, 128, 96), (1, 60, 16, 1, 128, 96), (3, 80, 16, 4, 128, 96), (2, 48, 8, 2, 64, 64)]
    for i, (B, N_CTX, Hq, Hkv, D, Dv) in enumerate(cases):
        pre = torch.randint(1, N_CTX // 2, (B,), dtype=torch.int64)
        ext = torch.randint(1, N_CTX // 2, (B,), dtype=torch.int64)
        sl = pre + ext
        start = torch.zeros(B, dtype=torch.int64)
        start[1:] = torch.cumsum(sl[:-1], 0)
        ext_start = torch.zeros(B, dtype=torch.int64)
        ext_start[1:] = torch.cumsum(ext[:-1], 0)
        T, Te = int(sl.sum()), int(ext.sum())
        r2t = torch.zeros(B, int(sl.max()), dtype=torch.int32)
        perm = torch.randperm(T).to(torch.int32)
        for b in range(B):
            r2t[b, : sl[b]] = perm[start[b]: start[b] + sl[b]]
        k_buf = torch.randn(T, Hkv, D, dtype=torch.bfloat16)
        v_buf = torch.randn(T, Hkv, Dv, dtype=torch.bfloat16)
        q = torch.randn(Te, Hq, D, dtype=torch.bfloat16)
        k_ext = torch.empty(Te, Hkv, D, dtype=torch.bfloat16)
        v_ext = torch.empty(Te, Hkv, Dv, dtype=torch.bfloat16)
        for b in range(B):
            slots = r2t[b, pre[b]: sl[b]].long()
            k_ext[ext_start[b]: ext_start[b] + ext[b]] = k_buf[slots]
            v_ext[ext_start[b]: ext_start[b] + ext[b]] = v_buf[slots]
        o = torch.empty(Te, Hq, Dv, dtype=torch.bfloat16)
        # extend.cpp requires extend_seq_lens / extend_start_loc in req_to_token's dtype, seq_lens int64
        ref.extend_attention_cpu(q, k_ext, v_ext, o, k_buf, v_buf, r2t, torch.arange(B, dtype=torch.int64), sl,
                                 ext.to(torch.int32), ext_start.to(torch.int32), int(ext.max()), 1.0 / D ** 0.5, 0.0)
        p = f"c{i}_"
        out.update({p + "shape": np.array([B, Hq, Hkv, D, Dv]), p + "q": u16(q), p + "k_buf": u16(k_buf), p + "v_buf": u16(v_buf),
                    p + "r2t": r2t.numpy(), p + "seq_lens": sl.numpy(), p + "ext_lens": ext.numpy(),
                    p + "ext_start": ext_start.numpy(), p + "o": u16(o)})
    out["n_cases"] = np.array(len(cases))
    save("extend_attention_cpu", **out)


# ------------------------------------------------------------------------------------------------
def gen_gemm(ref):
    torch.manual_seed(1234)
    t_fp8 = load_ref_module("sgl-kernel/tests/test_fp8_blockwise_gemm.py", "ref_test_fp8_blockwise_gemm")
    t_i8 = load_ref_module("sgl-kernel/tests/test_int8_gemm.py", "ref_test_int8_gemm")
    t_awq = load_ref_module("sgl-kernel/tests/test_awq_dequant.py", "ref_test_awq_dequant")
    t_blk = load_ref_module("python/sglang/test/test_block_fp8.py", "ref_test_block_fp8")
    cpu_utils = load_ref_module("test/srt/cpu/utils.py", "ref_cpu_utils")
    out = {}

    # --- fp8_blockwise_scaled_mm vs baseline_scaled_mm (test_fp8_blockwise_gemm.py:62-81) ---
    shapes = [(1, 128, 512), (5, 512, 1024), (127, 128, 512), (16, 384, 640), (33, 256, 256)]
    for i, (M, N, K) in enumerate(shapes):
        a = ((torch.rand(M, K) - 0.5) * 2 * 448.0).clamp(-448, 448).to(torch.float8_e4m3fn)
        b = ((torch.rand(N, K) - 0.5) * 2 * 448.0).clamp(-448, 448).to(torch.float8_e4m3fn).t()
        sa = (torch.randn(t_fp8.scale_shape(a.shape, (1, 128))) * 0.001).t().contiguous().t()
        sb = (torch.randn(t_fp8.scale_shape(b.shape, (128, 128))) * 0.001).t().contiguous().t()
        o = t_fp8.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
        # second reference form: native_w8a8_block_fp8_matmul(A, W[N,K], As, Ws[N/128,K/128]) (test_block_fp8.py:273)
        o2 = t_blk.native_w8a8_block_fp8_matmul(a, b.t().contiguous(), sa.contiguous(), sb.t().contiguous(), [128, 128],
                                                torch.bfloat16)
        p = f"fp8_{i}_"
        out.update({p + "a": u8(a), p + "w": u8(b.t()), p + "sa": sa.contiguous().numpy(), p + "sb": sb.contiguous().numpy(),
                    p + "o": u16(o), p + "o_native": u16(o2)})
    out["fp8_n"] = np.array(len(shapes))

    # --- per-token-group fp8 activation quant (test_block_fp8.py:23-48) ---
    x = (torch.randn(7, 512) * 3).to(torch.bfloat16)
    x[3, 128:256] = 0  # an all-zero group exercises eps
    xq, xs = t_blk.native_per_token_group_quant_fp8(x, 128)
    out.update({"gq_x": u16(x), "gq_q": u8(xq), "gq_s": xs.numpy()})

    # --- int8_scaled_mm vs torch_scaled_mm (test_int8_gemm.py:10-30) ---
    shapes = [(1, 16, 512), (16, 128, 512), (32, 136, 1024), (130, 256, 512)]
    for i, (M, N, K) in enumerate(shapes):
        a = t_i8.to_int8(torch.randn(M, K) * 5)
        b = t_i8.to_int8(torch.randn(N, K).t() * 5)
        sa, sb = torch.randn(M), torch.randn(N)
        bias = torch.randn(N).to(torch.bfloat16) * 10 if i % 2 == 0 else None
        o = t_i8.torch_scaled_mm(a, b, sa, sb, torch.bfloat16, bias)
        p = f"i8_{i}_"
        out.update({p + "a": a.numpy(), p + "w": b.t().contiguous().numpy(), p + "sa": sa.numpy(), p + "sb": sb.numpy(),
                    p + "o": u16(o)})
        if bias is not None:
            out[p + "bias"] = u16(bias)
    out["i8_n"] = np.array(len(shapes))

    # --- per_token_quant_int8 (test/srt/cpu/utils.py) and the reference CPU kernel per_token_quant_int8_cpu ---
    x = (torch.randn(9, 384) * 2).to(torch.bfloat16)
    q, s = cpu_utils.per_token_quant_int8(x)
    q_cpu, s_cpu = ref.per_token_quant_int8_cpu(x)  # uint8 = q + 128 (gemm_int8.cpp:388-393)
    out.update({"tq_x": u16(x), "tq_q": q.numpy(), "tq_s": s.numpy(), "tq_q_cpu_u8": q_cpu.numpy(), "tq_s_cpu": s_cpu.numpy()})

    # --- int8 W8A8 linear: int8_scaled_mm_with_quant (reference CPU kernel; test/srt/cpu/test_gemm.py:118-150) ---
    M, N, K = 6, 64, 256
    x = (torch.randn(M, K) / (K ** 0.5)).to(torch.bfloat16)
    w = (torch.randn(N, K) * 5).round().clamp(-127, 127).to(torch.int8)
    ws = torch.rand(N) * 0.01
    bias = torch.randn(N) * 0.1
    o = ref.int8_scaled_mm_with_quant(x, w, ws, bias, torch.bfloat16, False)
    out.update({"i8l_x": u16(x), "i8l_w": w.numpy(), "i8l_ws": ws.numpy(), "i8l_bias": bias.numpy(), "i8l_o": u16(o)})

    # --- fp8 W8A16 linear: fp8_scaled_mm_cpu (reference CPU kernel; test/srt/cpu/test_gemm.py:152-185) ---
    M, N, K = 5, 256, 384
    x = (torch.randn(M, K) / (K ** 0.5)).to(torch.bfloat16)
    wq, wscale, _ = cpu_utils.convert_weight(torch.randn(N, K) * 0.05, [128, 128], torch.bfloat16)
    o = ref.fp8_scaled_mm_cpu(x, wq, wscale.float().contiguous(), [128, 128], None, torch.bfloat16, False)
    out.update({"f8l_x": u16(x), "f8l_w": u8(wq), "f8l_ws": wscale.float().numpy(), "f8l_o": u16(o)})

    # --- awq_dequantize vs awq_dequantize_torch (test_awq_dequant.py:28-57; group = 128 like the models) ---
    for i, (K, C, dt) in enumerate([(256, 16, torch.float16), (384, 72, torch.bfloat16), (128, 32, torch.bfloat16)]):
        qw = torch.randint(0, torch.iinfo(torch.int32).max, (K, C), dtype=torch.int32)
        qz = torch.randint(0, torch.iinfo(torch.int32).max, (K // 128, C), dtype=torch.int32)
        sc = torch.rand(K // 128, C * 8).to(dt)
        o = t_awq.awq_dequantize_torch(qw, sc, qz, 128)
        assert o.dtype == dt
        p = f"awq_{i}_"
        out.update({p + "qw": qw.numpy(), p + "qz": qz.numpy(), p + "sc": u16(sc), p + "o": u16(o), p + "bf16": np.array(dt == torch.bfloat16)})
    out["awq_n"] = np.array(3)
    save("quant_gemm", **out)


def gen_indices():
    """test/srt/test_create_kvindices.py:18-64 (reference expression: torch.cat of req_to_token rows) and the
    page-table form of utils.py:44-94 evaluated with the same expression."""
    rng = np.random.RandomState(1234)
    max_batch, max_ctx = 64, 257
    r2t = torch.from_numpy(rng.permutation(max_batch * max_ctx).astype(np.int32)).reshape(max_batch, max_ctx)
    out = {"r2t": r2t.numpy()}
    for i, batch in enumerate([1, 7, 37]):
        rpi = torch.from_numpy(rng.choice(range(max_batch), size=batch, replace=False)).to(torch.int32)
        lens = torch.from_numpy(rng.choice(range(max_ctx), size=batch, replace=False)).to(torch.int32)
        indptr = torch.zeros(batch + 1, dtype=torch.int32)
        indptr[1:] = torch.cumsum(lens, 0)
        ref_idx = torch.cat([r2t[rpi[j], : lens[j]] for j in range(batch)], 0).contiguous()
        start = torch.from_numpy(rng.randint(0, 3, size=batch)).to(torch.int32)
        lens2 = torch.clamp(lens - start, min=0)
        indptr2 = torch.zeros(batch + 1, dtype=torch.int32)
        indptr2[1:] = torch.cumsum(lens2, 0)
        ref_idx2 = torch.cat([r2t[rpi[j], start[j]: start[j] + lens2[j]] for j in range(batch)], 0).contiguous()
        P = 16
        max_pages = (max_ctx + P - 1) // P
        pt = torch.full((batch, max_pages), -1, dtype=torch.int32)
        for j in range(batch):
            n = (int(lens[j]) + P - 1) // P
            pt[j, :n] = r2t[rpi[j], 0: n * P: P][:n] // P
        p = f"c{i}_"
        out.update({p + "rpi": rpi.numpy(), p + "lens": lens.numpy(), p + "indptr": indptr.numpy(), p + "kv_indices": ref_idx.numpy(),
                    p + "start": start.numpy(), p + "lens2": lens2.numpy(), p + "indptr2": indptr2.numpy(),
                    p + "kv_indices2": ref_idx2.numpy(), p + "page_table": pt.numpy()})
    out["n_cases"] = np.array(3)
    out["page_size"] = np.array(16)
    save("kv_indices", **out)


def main():
    assert os.path.isdir(REF), f"{REF} is not mounted; the fixtures are generated in the authoring container only"
    from oracle import ref_ops

    assert ref_ops.available(), ref_ops.why_unavailable()
    ref = ref_ops.ops()
    torch.set_num_threads(4)
    gen_decode(ref)
    gen_extend(ref)
    gen_gemm(ref)
    gen_indices()


if __name__ == "__main__":
    main()
