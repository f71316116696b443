"""CPU tier: the plain-C oracle (oracle/oracle.c) against the golden vectors produced by the reference itself
(tests/golden/make_golden.py: the reference's CPU kernels compiled from source + its Python test helpers).
This is what pins the oracle; the GPU tests then compare the CUDA path with the oracle."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as orc
from tests.util import bf16, f16, fp8, golden, golden_decode_cases, golden_extend_cases, max_abs, rel_l2


@pytest.mark.parametrize("case", list(golden_decode_cases()), ids=lambda c: f"c{c['idx']}_B{c['B']}H{c['Hq']}_{c['Hkv']}D{c['D']}_{c['Dv']}")
def test_decode_attention_matches_reference_cpu_kernel(case):
    c = case
    kb = c["k_buf"].clone()
    vb = kb.narrow(2, 0, c["Dv"]) if c["mla"] else c["v_buf"].clone()
    out = torch.zeros(c["B"], c["Hq"], c["Dv"], dtype=torch.bfloat16)
    # quirk_empty_split=True reproduces the reference literally: a kv split that holds no tokens (only possible
    # for seq_len < 57) still enters its merge with lse = 0 from the zero-initialised attn_logits (oracle.c)
    orc.decode_attention(c["q"], kb, vb, out, c["key"], c["value"], c["loc"], None, c["r2t"], c["rpi"], c["seq_lens"],
                         1.0 / c["D"] ** 0.5, c["cap"], num_kv_splits=8, is_mla=c["mla"], quirk_empty_split=True)
    # same algorithm (8 fixed splits, fp32 accumulate, MLA rounds P to bf16): differences are summation-order only
    assert rel_l2(out, c["o"]) < 2e-3
    assert max_abs(out, c["o"]) <= 2 ** -6
    # the contract the CUDA path is held to (exact softmax, empty splits skipped) equals the reference whenever
    # no split is empty
    split = (c["seq_lens"] + 7) // 8
    if bool(((c["seq_lens"] + split - 1) // split == 8).all()):
        out2 = torch.zeros_like(out)
        orc.decode_attention(c["q"], c["k_buf"].clone(), vb, out2, c["key"], c["value"], c["loc"], None, c["r2t"], c["rpi"],
                             c["seq_lens"], 1.0 / c["D"] ** 0.5, c["cap"], num_kv_splits=8, is_mla=c["mla"])
        assert torch.equal(out, out2)
    # the fused KV write is a bit-exact scatter (decode.cpp:771-809)
    exp_k = c["k_buf"].clone()
    exp_k[c["loc"]] = c["key"]
    assert torch.equal(kb, exp_k)
    if not c["mla"]:
        exp_v = c["v_buf"].clone()
        exp_v[c["loc"]] = c["value"]
        assert torch.equal(vb, exp_v)


@pytest.mark.parametrize("case", list(golden_extend_cases()), ids=lambda c: f"c{c['idx']}")
def test_extend_attention_matches_reference_cpu_kernel(case):
    c = case
    out = torch.zeros_like(c["o"])
    orc.extend_attention(c["q"], c["k_ext"], c["v_ext"], out, c["k_buf"], c["v_buf"], c["r2t"], c["rpi"], c["seq_lens"],
                         c["ext_lens"], c["ext_start"], int(c["ext_lens"].max()), 1.0 / c["D"] ** 0.5, 0.0)
    # test/srt/cpu/test_extend.py:183 uses atol=rtol=1e-2 against SDPA; kernel-vs-restatement is tighter
    assert rel_l2(out, c["o"]) < 5e-3
    torch.testing.assert_close(out.float(), c["o"].float(), atol=1e-2, rtol=1e-2)


def test_fp8_blockwise_scaled_mm_matches_reference_helpers():
    g = golden("quant_gemm")
    for i in range(int(g["fp8_n"])):
        p = f"fp8_{i}_"
        a, w = fp8(g[p + "a"]), fp8(g[p + "w"])
        sa = torch.from_numpy(g[p + "sa"]).t().contiguous().t()   # M-major (fp8_blockwise_gemm_kernel.cu:374)
        sb = torch.from_numpy(g[p + "sb"]).t().contiguous().t()
        out = orc.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
        for name in ("o", "o_native"):
            ref = bf16(g[p + name])
            assert rel_l2(out, ref) < 1e-3, (i, name, rel_l2(out, ref))
            torch.testing.assert_close(out.float(), ref.float(), rtol=0.02, atol=1)  # test_fp8_blockwise_gemm.py:80-82


def test_per_token_group_quant_fp8_matches_reference_helper():
    g = golden("quant_gemm")
    x = bf16(g["gq_x"])
    q, s = orc.per_token_group_quant_fp8(x, 128, column_major_scales=False)
    ref_s = torch.from_numpy(g["gq_s"])
    # the torch helper clamps the bf16 absmax with eps (1e-10 rounds to 1.0005e-10 in bf16); the CUDA kernel the
    # product is pinned to clamps in fp32 (per_token_group_quant_8bit.cu:60-87) -> only all-zero groups differ
    zero_group = x.view(7, 4, 128).abs().amax(-1) == 0
    assert torch.equal(s[~zero_group], ref_s[~zero_group])
    assert torch.allclose(s[zero_group], torch.full_like(s[zero_group], 1e-10 / 448.0), rtol=1e-6)
    assert torch.allclose(s[zero_group], ref_s[zero_group], rtol=1e-3)
    assert np.array_equal(q.view(torch.uint8).numpy(), g["gq_q"])  # bit-exact e4m3 codes


def test_int8_scaled_mm_matches_reference_helper():
    g = golden("quant_gemm")
    for i in range(int(g["i8_n"])):
        p = f"i8_{i}_"
        a, w = torch.from_numpy(g[p + "a"]), torch.from_numpy(g[p + "w"])
        sa, sb = torch.from_numpy(g[p + "sa"]), torch.from_numpy(g[p + "sb"])
        bias = bf16(g[p + "bias"]) if p + "bias" in g.files else None
        out = orc.int8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, bias)
        ref = bf16(g[p + "o"])
        assert torch.equal(out, ref), (i, max_abs(out, ref))  # int32-exact accumulate, fp32 scale in the same order


def test_per_token_quant_int8_matches_reference():
    g = golden("quant_gemm")
    x = bf16(g["tq_x"])
    q, s = orc.per_token_quant_int8(x)
    assert np.array_equal(s.numpy(), g["tq_s"])
    # codes: identical except at exact .5 ties, where the last ulp of 127/absmax decides (torch evaluates
    # 127/absmax as 127 * (1/absmax), Triton's `/` is div.full, C and our CUDA kernel use the IEEE quotient)
    d0 = q.numpy().astype(np.int32) - g["tq_q"].astype(np.int32)
    exact = x.double() * (127.0 / x.double().abs().amax(-1, keepdim=True))
    tie = ((exact.abs() % 1.0) - 0.5).abs().numpy() < 1e-5
    assert np.abs(d0).max() <= 1 and not (d0 != 0)[~tie].any()
    # the reference CPU kernel stores q + 128 as uint8 (gemm_int8.cpp:388-393); rounding may differ by one code
    # where x*127/absmax is within an ulp of .5 (the CPU kernel multiplies by a precomputed 127/absmax)
    d = np.abs(q.numpy().astype(np.int32) + 128 - g["tq_q_cpu_u8"].astype(np.int32))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3
    assert np.allclose(s.numpy().reshape(-1), g["tq_s_cpu"].reshape(-1), rtol=1e-6)


def test_int8_linear_matches_reference_cpu_kernel():
    g = golden("quant_gemm")
    x, w = bf16(g["i8l_x"]), torch.from_numpy(g["i8l_w"])
    ws, bias = torch.from_numpy(g["i8l_ws"]), torch.from_numpy(g["i8l_bias"])
    xq, xs = orc.per_token_quant_int8(x)
    out = orc.int8_scaled_mm(xq, w.t(), xs, ws, torch.float32, None) if False else None
    acc = orc.int8_scaled_mm(xq, w.t(), xs, ws, torch.bfloat16, None).float() + bias
    ref = bf16(g["i8l_o"]).float()
    assert rel_l2(acc, ref) < 1e-2


def test_fp8_w8a16_linear_matches_reference_cpu_kernel():
    g = golden("quant_gemm")
    x, w, ws = bf16(g["f8l_x"]), fp8(g["f8l_w"]), torch.from_numpy(g["f8l_ws"])
    out = orc.fp8_w8a16_mm(x, w, ws, (128, 128), None, torch.bfloat16)
    assert rel_l2(out, bf16(g["f8l_o"])) < 1e-2


def test_awq_dequantize_matches_reference_helper():
    g = golden("quant_gemm")
    for i in range(int(g["awq_n"])):
        p = f"awq_{i}_"
        is_bf16 = bool(g[p + "bf16"])
        sc = bf16(g[p + "sc"]) if is_bf16 else f16(g[p + "sc"])
        out = orc.awq_dequantize(torch.from_numpy(g[p + "qw"]), sc, torch.from_numpy(g[p + "qz"]))
        ref = bf16(g[p + "o"]) if is_bf16 else f16(g[p + "o"])
        # awq_dequantize_torch: int8 (w - z) times a half-precision scale, one rounding == (w-z)*s in half
        assert torch.equal(out.view(torch.int16), ref.view(torch.int16)), i


def test_kv_indices_match_reference_expression():
    g = golden("kv_indices")
    r2t = torch.from_numpy(g["r2t"])
    P = int(g["page_size"])
    for i in range(int(g["n_cases"])):
        p = f"c{i}_"
        rpi, lens, indptr = (torch.from_numpy(g[p + n]) for n in ("rpi", "lens", "indptr"))
        out = orc.create_kv_indices(r2t, rpi, lens, indptr)
        assert np.array_equal(out.numpy(), g[p + "kv_indices"])
        out2 = orc.create_kv_indices(r2t, rpi, torch.from_numpy(g[p + "lens2"]), torch.from_numpy(g[p + "indptr2"]),
                                     torch.from_numpy(g[p + "start"]))
        assert np.array_equal(out2.numpy(), g[p + "kv_indices2"])
        pt = orc.create_page_table(r2t, rpi, lens, P, g[p + "page_table"].shape[1])
        assert np.array_equal(pt.numpy(), g[p + "page_table"])
        # int64 page table (decode.cpp:1441-1445 accepts both)
        assert np.array_equal(orc.create_kv_indices(r2t.long(), rpi, lens, indptr).numpy(), g[p + "kv_indices"])
