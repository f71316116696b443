"""GPU tier: quantised linears (fp8_blockwise_scaled_mm, int8_scaled_mm, awq_dequantize / W4A16 GEMM) and the
activation quantisers against the CPU oracle and the reference's golden outputs.  Tolerance (north_star):
quantised GEMM within 1e-2; integer paths (int32 accumulation, nibble extraction, fp8 codes) bit-exact."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as orc
from tests.util import bf16, f16, fp8, golden, make_fp8_gemm_case, max_abs, rel_l2

pytestmark = pytest.mark.gpu


def _col_major(t):  # [K,N] view with stride(0) == 1 on the device
    return t.t().contiguous().cuda().t()


def _run_fp8(a, b, sa, sb, dtype=torch.bfloat16):
    from iaas_sglang_b200 import ops

    out = ops.fp8_blockwise_scaled_mm(a.cuda(), _col_major(b), _col_major(sa), _col_major(sb), dtype)
    torch.cuda.synchronize()
    return out.cpu()


def test_fp8_blockwise_against_reference_golden():
    g = golden("quant_gemm")
    for i in range(int(g["fp8_n"])):
        p = f"fp8_{i}_"
        a, w = fp8(g[p + "a"]), fp8(g[p + "w"])
        sa, sb = torch.from_numpy(g[p + "sa"]), torch.from_numpy(g[p + "sb"])
        out = _run_fp8(a, w.t(), sa, sb)
        ref = bf16(g[p + "o"])
        assert rel_l2(out, ref) < 1e-2
        torch.testing.assert_close(out.float(), ref.float(), rtol=0.02, atol=1)  # test_fp8_blockwise_gemm.py:80-82


FP8_SHAPES = [(1, 128, 512), (3, 512, 1024), (5, 1024, 4096), (127, 512, 512), (128, 4096, 512), (512, 1024, 1024),
              (64, 6144, 4096), (64, 4096, 4096), (64, 4096, 14336), (256, 1280, 8192), (256, 8192, 1024),
              (300, 384, 640), (16, 14080, 512), (1024, 2048, 2048)]


@pytest.mark.parametrize("M,N,K", FP8_SHAPES)
@pytest.mark.parametrize("dist", ["uniform448", "realistic"])
def test_fp8_blockwise_against_oracle(M, N, K, dist):
    if dist == "realistic" and M * N * K > 64 * 6144 * 4096:
        pytest.skip("one distribution is enough for the largest shapes")
    a, b, sa, sb = make_fp8_gemm_case(M, N, K, seed=M + N + K, dist=dist)
    ref = orc.fp8_blockwise_scaled_mm(a, b, sa, sb, torch.bfloat16)
    out = _run_fp8(a, b, sa, sb)
    assert rel_l2(out, ref) < 1e-2, rel_l2(out, ref)
    assert rel_l2(out, ref) < 1e-3  # in practice: fp32 promotion per 128-block, bf16 output rounding only


# the TP-local shapes of SURVEY 8a row a11 at the decode batch sizes where the skinny / large-M crossover sits:
# cfg 4 (Llama-70B TP8), cfg 5 (DeepSeek TP8), cfg 2 under TP4, cfg 2's gate_up
TP_SHAPES = ([(m, n, k) for m in (129, 192, 256) for n, k in [(1280, 8192), (8192, 1024), (7168, 8192), (8192, 3584)]]
             + [(256, n, k) for n, k in [(2112, 7168), (3072, 1536), (7168, 2048), (4608, 7168), (7168, 2304)]]
             + [(256, n, k) for n, k in [(1536, 4096), (4096, 1024), (7168, 4096), (4096, 3584)]]
             + [(64, 28672, 4096), (128, 7168, 8192), (100, 2112, 7168)])


@pytest.mark.parametrize("M,N,K", TP_SHAPES)
def test_fp8_blockwise_tp_shapes_on_both_kernels(M, N, K):
    """Both FP8 kernels must be right on every shape either side of the crossover (knob sk_max_m): the measured
    policy may move it."""
    from iaas_sglang_b200 import _lib

    a, b, sa, sb = make_fp8_gemm_case(M, N, K, seed=M + N + K, dist="realistic")
    ref = orc.fp8_blockwise_scaled_mm(a, b, sa, sb, torch.bfloat16)
    old = _lib.tune_get("sk_max_m")
    try:
        for kernel, max_m in (("skinny", 256), ("large", 0), ("policy", old)):
            _lib.tune_set("sk_max_m", max_m)
            out = _run_fp8(a, b, sa, sb)
            assert rel_l2(out, ref) < 1e-3, (kernel, rel_l2(out, ref))
    finally:
        _lib.tune_set("sk_max_m", old)


@pytest.mark.parametrize("M,N,K,split", [(200, 1280, 8192, 0), (256, 1288, 8192, 7), (130, 2112, 7168, 4), (512, 768, 4096, 6),
                                         (129, 520, 2048, 3), (256, 4096, 1024, 2)])
def test_fp8_large_m_kernel_split_k(M, N, K, split):
    """The large-M kernel's split-K (few-tile shapes: K slices meet in the zero-initialised workspace behind self-resetting
    counters): right against the oracle, bit-reproducible from call to call (fixed summation order), and the counters are
    back at zero afterwards (the unsplit kernel and other shapes share the workspace).  Ragged M and N, forced and automatic
    slice counts (knob gb_split)."""
    from iaas_sglang_b200 import _lib

    a, b, sa, sb = make_fp8_gemm_case(M, N, K, seed=M + N + K, dist="realistic")
    ref = orc.fp8_blockwise_scaled_mm(a, b, sa, sb, torch.bfloat16)
    old_m, old_s = _lib.tune_get("sk_max_m"), _lib.tune_get("gb_split")
    try:
        _lib.tune_set("sk_max_m", 0)      # large-M kernel whatever M
        _lib.tune_set("gb_split", 1)
        plain = _run_fp8(a, b, sa, sb)
        _lib.tune_set("gb_split", split)
        outs = [_run_fp8(a, b, sa, sb) for _ in range(3)]
        torch.cuda.synchronize()
        assert rel_l2(outs[0], ref) < 1e-3, rel_l2(outs[0], ref)
        assert rel_l2(outs[0], plain) < 1e-3
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
        _lib.tune_set("gb_split", 1)
        assert torch.equal(_run_fp8(a, b, sa, sb), plain)
    finally:
        _lib.tune_set("sk_max_m", old_m)
        _lib.tune_set("gb_split", old_s)


@pytest.mark.parametrize("M,Hq,Hkv,K", [(64, 32, 8, 4096), (1, 32, 8, 4096), (37, 16, 4, 4096), (128, 24, 8, 4096),
                                        (64, 8, 2, 512), (100, 10, 2, 5120), (256, 12, 2, 4096)])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
def test_fp8_qkv_gemm_with_fused_rope_is_bit_identical_to_gemm_then_rope(M, Hq, Hkv, K, dt):
    """qkv projection + neox RoPE in the GEMM's epilogue (decode-shape kernel, head_dim 128) against the GEMM followed by
    ops.rope_neox_: the same bits on the q and k heads, v untouched; shapes the large-M kernel serves (M = 256) take the
    two-launch path inside the same call."""
    from iaas_sglang_b200 import ops

    D = 128
    N = (Hq + 2 * Hkv) * D
    a, b, sa, sb = make_fp8_gemm_case(M, N, K, seed=M + N + K, dist="realistic")
    g = torch.Generator().manual_seed(M)
    pos = torch.randint(0, 131072, (M,), generator=g, dtype=torch.int64).cuda()
    a, b, sa, sb = a.cuda(), _col_major(b), _col_major(sa), _col_major(sb)
    want = ops.fp8_blockwise_scaled_mm(a, b, sa, sb, dt)
    plain = want.clone()
    ops.rope_neox_(want[:, :Hq * D], want[:, Hq * D:(Hq + Hkv) * D], pos, D, 500000.0)
    cs = ops.rope_cos_sin(pos, D, 500000.0)
    got = ops.fp8_blockwise_scaled_mm(a, b, sa, sb, dt, rope=(pos, (Hq + Hkv) * D, D, 500000.0, cs))
    got2 = ops.fp8_blockwise_scaled_mm(a, b, sa, sb, dt, rope=(pos, (Hq + Hkv) * D, D, 500000.0, None))   # two launches
    assert torch.equal(got2.view(torch.int16), want.view(torch.int16))
    torch.cuda.synchronize()
    assert torch.equal(got[:, (Hq + Hkv) * D:], plain[:, (Hq + Hkv) * D:])        # v: plain GEMM output
    assert not torch.equal(got[:, :Hq * D], plain[:, :Hq * D]) or M == 0          # q was rotated
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


def test_fp8_blockwise_fp16_out_and_preconditions():
    from iaas_sglang_b200 import ops

    a, b, sa, sb = make_fp8_gemm_case(33, 256, 512, seed=3, dist="realistic")
    ref = orc.fp8_blockwise_scaled_mm(a, b, sa, sb, torch.float16)
    assert rel_l2(_run_fp8(a, b, sa, sb, torch.float16), ref) < 1e-3
    ad, bd, sad, sbd = a.cuda(), _col_major(b), _col_major(sa), _col_major(sb)
    with pytest.raises(RuntimeError, match="column major"):
        ops.fp8_blockwise_scaled_mm(ad, bd.contiguous(), sad, sbd, torch.bfloat16)
    with pytest.raises(RuntimeError, match="M major"):
        ops.fp8_blockwise_scaled_mm(ad, bd, sad.contiguous(), sbd, torch.bfloat16)
    with pytest.raises(RuntimeError, match="Half or BFloat16"):
        ops.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.float32)
    assert ops.fp8_blockwise_scaled_mm(ad[:0], bd, sad[:0], sbd, torch.bfloat16).shape == (0, 256)


def test_fp8_linearity_property_large():
    """Size-independent property at a prefill-size GEMM: scaling sa by 2 doubles the output bit-exactly; a sampled
    set of rows matches the oracle."""
    M, N, K = 4096, 4096, 4096
    a, b, sa, sb = make_fp8_gemm_case(M, N, K, seed=11, dist="realistic")
    out = _run_fp8(a, b, sa, sb)
    out2 = _run_fp8(a, b, sa * 2, sb)
    assert torch.equal(out2, out * 2)
    rows = torch.tensor([0, 1, 127, 128, 2049, 4095])
    ref = orc.fp8_blockwise_scaled_mm(a[rows], b, sa[rows].t().contiguous().t(), sb, torch.bfloat16)
    assert rel_l2(out[rows], ref) < 1e-3


def test_per_token_group_quant_fp8_bit_exact():
    from iaas_sglang_b200 import ops

    g = golden("quant_gemm")
    x = bf16(g["gq_x"])
    q, s = ops.per_token_group_quant_fp8(x.cuda(), 128)
    nz = x.view(7, 4, 128).abs().amax(-1) != 0
    assert torch.equal(s.cpu()[nz], torch.from_numpy(g["gq_s"])[nz])
    assert np.array_equal(q.cpu().view(torch.uint8).numpy(), g["gq_q"])
    for M, K, dt in [(64, 4096, torch.bfloat16), (1, 128, torch.float16), (255, 14336, torch.bfloat16), (3, 7168, torch.bfloat16)]:
        x = (torch.randn(M, K) * 3).to(dt)
        q0, s0 = orc.per_token_group_quant_fp8(x)
        for col_major in (True, False):
            q1, s1 = ops.per_token_group_quant_fp8(x.cuda(), 128, column_major_scales=col_major)
            assert s1.stride(0) == (1 if col_major and M > 1 else s1.stride(0))
            assert torch.equal(q0.view(torch.uint8), q1.cpu().view(torch.uint8)) and torch.equal(s0, s1.cpu())


def test_per_token_quant_int8_bit_exact():
    from iaas_sglang_b200 import ops

    for M, K in [(9, 384), (64, 4096), (1, 16), (130, 5120)]:
        x = (torch.randn(M, K) * 2).to(torch.bfloat16)
        q0, s0 = orc.per_token_quant_int8(x)
        q1, s1 = ops.per_token_quant_int8(x.cuda())
        assert torch.equal(q0, q1.cpu()) and torch.equal(s0, s1.cpu())


INT8_SHAPES = [(1, 16, 512), (16, 128, 512), (32, 136, 1024), (130, 256, 512), (64, 4096, 4096), (64, 6144, 4096),
               (512, 1024, 8192), (300, 136, 1024), (1024, 4096, 512)]


@pytest.mark.parametrize("M,N,K", INT8_SHAPES)
@pytest.mark.parametrize("with_bias", [True, False])
def test_int8_scaled_mm_bit_exact(M, N, K, with_bias):
    """test_int8_gemm.py:19-30 inputs; int32 accumulation is exact and the epilogue applies the scales in the
    reference's order, so the bf16 output equals the oracle bit for bit."""
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(M * 7 + N + K)
    a = (torch.randn(M, K, generator=g) * 5).clamp(-128, 127).round().to(torch.int8)
    w = (torch.randn(N, K, generator=g) * 5).clamp(-128, 127).round().to(torch.int8)
    sa, sb = torch.randn(M, generator=g), torch.randn(N, generator=g)
    bias = (torch.randn(N, generator=g) * 10).to(torch.bfloat16) if with_bias else None
    ref = orc.int8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, bias)
    out = ops.int8_scaled_mm(a.cuda(), w.cuda().t(), sa.cuda(), sb.cuda(), torch.bfloat16, bias.cuda() if with_bias else None)
    assert torch.equal(out.cpu(), ref), max_abs(out, ref)


@pytest.mark.parametrize("M,N,K", [(130, 1536, 4096), (200, 4096, 1024), (256, 1280, 8192), (256, 3072, 1536), (96, 4096, 512)])
def test_int8_scaled_mm_both_kernels_bit_exact(M, N, K):
    """Either side of the measured int8 policy (knob sk_max_m): the decode-shape kernel up to M = 256 and the large-M kernel
    from M = 1 both equal the oracle bit for bit, so the policy can move without changing a result."""
    from iaas_sglang_b200 import _lib, ops

    g = torch.Generator().manual_seed(M + N + K)
    a = (torch.randn(M, K, generator=g) * 5).clamp(-128, 127).round().to(torch.int8)
    w = (torch.randn(N, K, generator=g) * 5).clamp(-128, 127).round().to(torch.int8)
    sa, sb = torch.randn(M, generator=g), torch.randn(N, generator=g)
    bias = (torch.randn(N, generator=g) * 10).to(torch.bfloat16)
    ref = orc.int8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, bias)
    old = _lib.tune_get("sk_max_m")
    try:
        for kernel, max_m in (("skinny", 256), ("large", 0), ("policy", old)):
            _lib.tune_set("sk_max_m", max_m)
            out = ops.int8_scaled_mm(a.cuda(), w.cuda().t(), sa.cuda(), sb.cuda(), torch.bfloat16, bias.cuda())
            assert torch.equal(out.cpu(), ref), (kernel, max_abs(out, ref))
    finally:
        _lib.tune_set("sk_max_m", old)


def test_int8_against_reference_golden_and_extremes():
    from iaas_sglang_b200 import ops

    g = golden("quant_gemm")
    for i in range(int(g["i8_n"])):
        p = f"i8_{i}_"
        a, w = torch.from_numpy(g[p + "a"]), torch.from_numpy(g[p + "w"])
        bias = bf16(g[p + "bias"]).cuda() if p + "bias" in g.files else None
        out = ops.int8_scaled_mm(a.cuda(), w.cuda().t(), torch.from_numpy(g[p + "sa"]).cuda(),
                                 torch.from_numpy(g[p + "sb"]).cuda(), torch.bfloat16, bias)
        assert torch.equal(out.cpu(), bf16(g[p + "o"]))
    # saturating inputs: -128 * -128 * K stays inside int32 for K = 16384
    a = torch.full((4, 16384), -128, dtype=torch.int8)
    w = torch.full((16, 16384), -128, dtype=torch.int8)
    one_m, one_n = torch.ones(4), torch.ones(16)
    out = ops.int8_scaled_mm(a.cuda(), w.cuda().t(), one_m.cuda(), one_n.cuda(), torch.bfloat16)
    assert torch.equal(out.cpu(), orc.int8_scaled_mm(a, w.t(), one_m, one_n, torch.bfloat16))


@pytest.mark.parametrize("K,C,dt", [(256, 16, torch.float16), (384, 72, torch.bfloat16), (3584, 448, torch.bfloat16),
                                    (5120, 448, torch.float16), (128, 32, torch.bfloat16)])
def test_awq_dequantize_bit_exact(K, C, dt):
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(K + C)
    qw = torch.randint(0, 2 ** 31 - 1, (K, C), generator=g, dtype=torch.int32)
    qz = torch.randint(0, 2 ** 31 - 1, (K // 128, C), generator=g, dtype=torch.int32)
    sc = torch.rand(K // 128, C * 8, generator=g).to(dt)
    ref = orc.awq_dequantize(qw, sc, qz)
    out = ops.awq_dequantize(qw.cuda(), sc.cuda(), qz.cuda())
    assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16))


def test_awq_dequantize_reference_golden():
    from iaas_sglang_b200 import ops

    g = golden("quant_gemm")
    for i in range(int(g["awq_n"])):
        p = f"awq_{i}_"
        is_bf16 = bool(g[p + "bf16"])
        sc = bf16(g[p + "sc"]) if is_bf16 else f16(g[p + "sc"])
        out = ops.awq_dequantize(torch.from_numpy(g[p + "qw"]).cuda(), sc.cuda(), torch.from_numpy(g[p + "qz"]).cuda())
        ref = bf16(g[p + "o"]) if is_bf16 else f16(g[p + "o"])
        assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16))


@pytest.mark.parametrize("M,K,N,dt", [(16, 512, 1024, torch.bfloat16), (64, 1024, 512, torch.float16),
                                      (128, 5120, 3584, torch.bfloat16), (1, 2560, 5120, torch.bfloat16),
                                      (130, 1024, 1088, torch.float16)])
def test_awq_w4a16_gemm(M, K, N, dt):
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(M + K + N)
    qw = torch.randint(0, 2 ** 31 - 1, (K, N // 8), generator=g, dtype=torch.int32)
    qz = torch.randint(0, 2 ** 31 - 1, (K // 128, N // 8), generator=g, dtype=torch.int32)
    sc = (torch.rand(K // 128, N, generator=g) * 0.01).to(dt)
    x = torch.randn(M, K, generator=g).to(dt)
    bias = torch.randn(N, generator=g).to(dt) if M % 2 == 0 else None
    ref = orc.awq_gemm(x, qw, sc, qz, bias)
    out = ops.awq_w4a16_gemm(x.cuda(), qw.cuda(), sc.cuda(), qz.cuda(), bias.cuda() if bias is not None else None)
    assert rel_l2(out, ref) < 1e-2
    assert rel_l2(out, ref) < 2e-3


@pytest.mark.parametrize("K,N,group,dt", [(128, 256, 128, torch.float16), (256, 1088, 128, torch.bfloat16),
                                          (256, 512, 32, torch.float16), (192, 160, 64, torch.bfloat16)])
def test_awq_fused_gemm_dequant_is_bit_exact(K, N, group, dt):
    """Identity activations read the dequantised weight back through the fused kernel (int4 box -> registers ->
    swizzled tcgen05 operand tile -> TMEM -> permuted write-out): it must equal awq_dequantize bit for bit."""
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(K + N + group)
    qw = torch.randint(-2 ** 31, 2 ** 31 - 1, (K, N // 8), generator=g, dtype=torch.int32)
    qz = torch.randint(-2 ** 31, 2 ** 31 - 1, (K // group, N // 8), generator=g, dtype=torch.int32)
    sc = (torch.rand(K // group, N, generator=g) * 0.02 + 1e-3).to(dt)
    ref = orc.awq_dequantize(qw, sc, qz)
    out = ops.awq_w4a16_gemm(torch.eye(K, dtype=dt).cuda(), qw.cuda(), sc.cuda(), qz.cuda())
    assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16))


@pytest.mark.parametrize("M", [1, 7, 64, 128, 200, 256])
@pytest.mark.parametrize("N,K", [(3584, 5120), (5120, 2560), (27648, 5120), (5120, 13824)])  # cfg 3 under TP2
def test_awq_fused_gemm_cfg3_shapes(M, N, K):
    from iaas_sglang_b200 import ops

    dt = torch.bfloat16
    g = torch.Generator().manual_seed(M + K + N)
    qw = torch.randint(-2 ** 31, 2 ** 31 - 1, (K, N // 8), generator=g, dtype=torch.int32)
    qz = torch.randint(-2 ** 31, 2 ** 31 - 1, (K // 128, N // 8), generator=g, dtype=torch.int32)
    sc = (torch.rand(K // 128, N, generator=g) * 0.01 + 1e-3).to(dt)
    x = torch.randn(M, K, generator=g).to(dt)
    bias = torch.randn(N, generator=g).to(dt) if M % 2 == 0 else None
    # awq.py:186-204 literally: dequantise (oracle, bit-exact to awq_dequantize), then matmul (fp32 accumulation)
    ref = x.float() @ orc.awq_dequantize(qw, sc, qz).float()
    if bias is not None:
        ref += bias.float()
    out = ops.awq_w4a16_gemm(x.cuda(), qw.cuda(), sc.cuda(), qz.cuda(), bias.cuda() if bias is not None else None)
    assert rel_l2(out.cpu().float(), ref) < 2e-3
    # the same call again gives the same bits (fixed-order split-K reduction)
    out2 = ops.awq_w4a16_gemm(x.cuda(), qw.cuda(), sc.cuda(), qz.cuda(), bias.cuda() if bias is not None else None)
    assert torch.equal(out, out2)


# ---- fused producers of an FP8 linear's input (fused_elem.cu) ----
@pytest.mark.parametrize("M,H,dt,res", [(64, 4096, torch.bfloat16, True), (7, 2048, torch.bfloat16, False),
                                        (130, 8192, torch.bfloat16, True), (33, 5120, torch.float16, True),
                                        (1, 128, torch.bfloat16, True)])
def test_rmsnorm_quant_fp8_is_bit_identical_to_the_two_step_path(M, H, dt, res):
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(M + H)
    x = (torch.randn(M, H, generator=g) * 2).to(dt).cuda()
    w = (torch.rand(H, generator=g) + 0.5).to(dt).cuda()
    r1 = (torch.randn(M, H, generator=g)).to(dt).cuda() if res else None
    r2 = r1.clone() if res else None
    normed = ops.rmsnorm(x, w, 1e-5, residual=r1)
    q_ref, s_ref = ops.per_token_group_quant_fp8(normed, 128, column_major_scales=True)
    q, s, half = ops.rmsnorm_quant_fp8(x, w, 1e-5, residual=r2, want_half=True)
    assert torch.equal(half.view(torch.int16), normed.view(torch.int16))
    assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8)) and torch.equal(s, s_ref)
    assert s.stride(0) == 1  # column-major scales, as fp8_blockwise_scaled_mm wants them
    if res:
        assert torch.equal(r1.view(torch.int16), r2.view(torch.int16))
    # and the norm itself against the oracle (layernorm.py forward_native)
    xs = x.float().cpu() + (r1.float().cpu() - x.float().cpu() if False else 0)
    base = r1.float().cpu() if res else x.float().cpu()   # r1 now holds x + residual
    ref = base * torch.rsqrt(base.pow(2).mean(-1, keepdim=True) + 1e-5)
    ref = ref.to(dt).float() * w.float().cpu()
    assert rel_l2(normed.float().cpu(), ref) < 3e-3


@pytest.mark.parametrize("M,d,dt", [(64, 14336, torch.bfloat16), (5, 128, torch.bfloat16), (129, 3584, torch.float16)])
def test_silu_mul_quant_fp8_is_bit_identical_to_the_two_step_path(M, d, dt):
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(M + d)
    x = (torch.randn(M, 2 * d, generator=g) * 3).to(dt).cuda()
    act = ops.silu_and_mul(x)
    q_ref, s_ref = ops.per_token_group_quant_fp8(act, 128, column_major_scales=True)
    q, s = ops.silu_and_mul_quant_fp8(x)
    assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8)) and torch.equal(s, s_ref)


def test_fp8_linear_with_prequantised_input_equals_apply():
    from iaas_sglang_b200 import linear, ops

    class L:  # the parameter names Fp8LinearMethod creates (fp8.py:236-300)
        pass

    g = torch.Generator().manual_seed(5)
    M, K, N = 64, 4096, 6144
    layer = L()
    layer.weight = (torch.randn(N, K, generator=g) * 30).clamp(-448, 448).to(torch.float8_e4m3fn).cuda()
    layer.weight_scale_inv = (torch.rand(N // 128, K // 128, generator=g) * 1e-2 + 1e-3).cuda()
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).cuda()
    w = torch.ones(K, dtype=torch.bfloat16).cuda()
    m = linear.Fp8BlockLinearMethod()
    a = m.apply(layer, ops.rmsnorm(x, w, 1e-5))
    xq, xs = ops.rmsnorm_quant_fp8(x, w, 1e-5)
    b = m.apply_quantized(layer, xq, xs, torch.bfloat16)
    assert torch.equal(a, b)


@pytest.mark.parametrize("Hq,Hkv,D", [(32, 8, 128), (10, 2, 128), (32, 8, 64), (3, 1, 80)])
def test_rope_neox_against_forward_native(Hq, Hkv, D):
    """rotary_embedding.py forward_native (is_neox_style): x1' = x1 cos - x2 sin, x2' = x2 cos + x1 sin."""
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(Hq + D)
    M = 37
    qkv = torch.randn(M, (Hq + 2 * Hkv) * D, generator=g).to(torch.bfloat16).cuda()
    q, k, _ = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    pos = torch.randint(0, 8192, (M,), generator=g, dtype=torch.int64).cuda()
    q0, k0 = q.float().cpu().view(M, Hq, D), k.float().cpu().view(M, Hkv, D)
    ops.rope_neox_(q, k, pos, D, 500000.0)
    inv = 1.0 / (500000.0 ** (torch.arange(0, D, 2).float() / D))
    ang = pos.cpu().float()[:, None] * inv[None]
    cs, sn = ang.cos()[:, None, :], ang.sin()[:, None, :]

    def rot(t):
        x1, x2 = t[..., : D // 2], t[..., D // 2:]
        return torch.cat([x1 * cs - x2 * sn, x2 * cs + x1 * sn], -1)

    assert rel_l2(q.float().cpu().view(M, Hq, D), rot(q0)) < 6e-3
    assert rel_l2(k.float().cpu().view(M, Hkv, D), rot(k0)) < 6e-3
