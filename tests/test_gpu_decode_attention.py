"""GPU tier: paged decode attention (b200_decode_attn through ops.decode_attention) against the CPU oracle and
the reference's golden outputs.  Tolerance (BASELINE.json north_star): 2e-3 relative error in bf16; the page-table
indexing and the fused KV append are bit-exact."""
import pytest
import torch

from tests.util import DecodeCase, golden_decode_cases, max_abs, rel_l2, run_cuda_decode

pytestmark = pytest.mark.gpu
TOL = 2e-3
# MLA only: the reference's CPU kernel rounds every softmax weight to bf16 before P.V (decode.cpp:1157-1159), which
# puts its OWN output ~2e-3 (relative L2, randn inputs) away from the exact softmax.  The CUDA kernel is held to
# 2e-3 against the exact form (fp32 weights) and to 2x that against the reference's literally-rounded output -- inside
# the bar the reference's own MLA kernel test grants (sgl-kernel/tests/test_cutlass_mla.py:99: atol = rtol = 1e-2), which
# tests/test_gpu_mla_boundary.py applies verbatim to the cutlass_mla_decode surface.
TOL_MLA_LITERAL = 4e-3


def check(c: DecodeCase, cap=0.0, fused=False, num_splits=0, write_kv=True, first_slot=0):
    ref, kb, vb = c.oracle(logit_cap=cap, write_kv=write_kv)
    out, kb2, vb2 = run_cuda_decode(c, logit_cap=cap, fused=fused, num_splits=num_splits, write_kv=write_kv)
    assert not torch.isnan(out.float()).any()
    err = rel_l2(out, ref)
    if c.mla:
        assert err < TOL_MLA_LITERAL, f"rel_l2 vs reference-rounded {err:.3e}"
        ref, _, _ = c.oracle(logit_cap=cap, write_kv=write_kv, round_p=False)
        err = rel_l2(out, ref)
    assert err < TOL, f"rel_l2 {err:.3e} maxabs {max_abs(out, ref):.3e}"
    # first_slot=1 skips the reserved dummy slot 0 (memory_pool.py:236-239): several padded rows write it at once,
    # so its content is whichever write lands last -- unspecified in the reference too
    assert torch.equal(kb[first_slot:], kb2[first_slot:]) and torch.equal(vb[first_slot:], vb2[first_slot:]), \
        "KV pool differs from the oracle's after the append"
    return out


# the reference's own shapes, test/srt/cpu/test_decode.py:146-158 (seq_len 1024, B 1-2) + test_mla.py:146-150
REF_SHAPES = [(2, 16, 16, 64, 64), (2, 16, 1, 16, 16), (2, 32, 8, 33, 55), (2, 16, 1, 64, 64), (2, 64, 1, 13, 13),
              (2, 128, 1, 80, 80), (2, 128, 2, 512, 512), (1, 16, 1, 576, 512), (1, 16, 16, 576, 512), (1, 22, 1, 576, 512),
              (1, 40, 8, 128, 128)]


@pytest.mark.parametrize("shape", REF_SHAPES, ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("fused", [False, True], ids=["append_kernel", "fused_append"])
def test_reference_test_shapes(shape, fused):
    B, Hq, Hkv, D, Dv = shape
    check(DecodeCase(B, Hq, Hkv, D, Dv, 1024), fused=fused)


@pytest.mark.parametrize("B,S", [(1, 8 * 111), (4, 8 * 128), (40, 8 * 133)])
def test_mla_reference_shapes(B, S):
    check(DecodeCase(B, 22, 1, 576, 512, S), fused=True)


@pytest.mark.parametrize("case", list(golden_decode_cases()), ids=lambda c: f"c{c['idx']}")
def test_against_reference_golden_outputs(case):
    """Outputs of the reference's decode_attention_cpu itself (tests/golden).  Requests whose 8 fixed splits are not
    all populated are excluded: there the reference merges an lse=0 phantom split (oracle.c quirk_empty_split)."""
    from iaas_sglang_b200 import ops

    c = case
    d = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in c.items()}
    kb = d["k_buf"]
    vb = kb.narrow(2, 0, c["Dv"]) if c["mla"] else d["v_buf"]
    key = d["key"]
    val = key.narrow(2, 0, c["Dv"]) if c["mla"] else d["value"]
    out = torch.zeros(c["B"], c["Hq"], c["Dv"], dtype=torch.bfloat16, device="cuda")
    ops.decode_attention(d["q"], kb, vb, out, key, val, d["loc"], None, d["r2t"], d["rpi"], d["seq_lens"],
                         1.0 / c["D"] ** 0.5, c["cap"], max_seq_len=c["S"])
    split = (c["seq_lens"] + 7) // 8
    full = (c["seq_lens"] + split - 1) // split == 8
    if full.any():
        assert rel_l2(out.cpu()[full], c["o"][full]) < (TOL_MLA_LITERAL if c["mla"] else TOL)
    exp_k = c["k_buf"].clone()
    exp_k[c["loc"]] = c["key"]
    assert torch.equal(kb.cpu(), exp_k)


@pytest.mark.parametrize("seq_lens", [[1], [1, 2, 3, 17], [128, 129, 127, 1], [5000, 3, 977], [4096] * 3])
@pytest.mark.parametrize("heads", [(32, 8, 128), (8, 1, 128), (20, 4, 128), (10, 2, 128), (32, 8, 64), (64, 2, 128)],
                         ids=lambda h: f"Hq{h[0]}Hkv{h[1]}D{h[2]}")
def test_ragged_lengths_and_head_layouts(seq_lens, heads):
    Hq, Hkv, D = heads
    check(DecodeCase(len(seq_lens), Hq, Hkv, D, D, seq_lens), fused=True)


@pytest.mark.parametrize("num_splits", [1, 2, 3, 8, 16, 64])
def test_split_counts(num_splits):
    check(DecodeCase(3, 32, 8, 128, 128, [2048, 100, 7]), num_splits=num_splits)
    check(DecodeCase(2, 16, 1, 576, 512, [600, 9]), num_splits=num_splits)


def test_logit_cap_and_fp16():
    check(DecodeCase(2, 32, 8, 128, 128, 2048), cap=30.0)
    check(DecodeCase(2, 32, 8, 64, 64, 300), cap=5.0, fused=True)
    check(DecodeCase(2, 32, 8, 33, 55, 300), cap=30.0)
    c = DecodeCase(2, 32, 8, 128, 128, 700, dtype=torch.float16)
    ref, _, _ = c.oracle()
    out, _, _ = run_cuda_decode(c)
    assert rel_l2(out, ref) < 5e-4  # fp16 has 3 more mantissa bits


def test_int64_page_table_and_scattered_request_slots():
    check(DecodeCase(4, 32, 8, 128, 128, [700, 1, 64, 2000], idx_dtype=torch.int64, extra_reqs=9), fused=True)
    check(DecodeCase(3, 16, 1, 576, 512, [70, 1, 640], idx_dtype=torch.int64, extra_reqs=5), fused=True)


def test_shared_prefix_radix_layout():
    """config 4's RadixAttention emulation: every request's first 256 positions point at the same slots."""
    check(DecodeCase(6, 8, 1, 128, 128, 900, shared_prefix=256), fused=True)
    check(DecodeCase(4, 32, 8, 128, 128, 600, shared_prefix=599), fused=True)


def test_head_major_non_contiguous_pool():
    check(DecodeCase(2, 32, 8, 128, 128, 513, noncontig=True))
    check(DecodeCase(2, 32, 8, 33, 55, 200, noncontig=True))


def test_no_kv_write_and_cuda_graph_padding_rows():
    check(DecodeCase(2, 32, 8, 128, 128, 300), write_kv=False)
    # cuda-graph padded rows: seq_len 1, req_pool_index 0, out_cache_loc 0 (triton_backend.py:629-630)
    c = DecodeCase(4, 32, 8, 128, 128, [300, 1, 1, 1])
    c.req_pool_indices[1:] = 0
    c.loc[1:] = 0
    check(c, first_slot=1)
    check(c, first_slot=1, fused=True)


def test_unique_slot_indexing_is_exact():
    """One-hot probe of the page-table walk: V row of slot s holds the value s (exactly representable), K = 0 so the
    softmax is uniform; the output must equal the mean of the request's slot ids."""
    from iaas_sglang_b200 import ops

    B, Hq, Hkv, D = 3, 8, 2, 128
    lens = torch.tensor([64, 256, 16], dtype=torch.int64)
    T = 2048
    g = torch.Generator().manual_seed(7)
    r2t = torch.zeros(B, 256, dtype=torch.int32)
    slots_all = torch.randperm(T - 1, generator=g)[: int(lens.sum())] + 1
    off = 0
    means = []
    for b in range(B):
        s = slots_all[off: off + int(lens[b])]
        off += int(lens[b])
        # restrict slots to < 256 multiples so the mean is exactly representable in bf16 after fp32 accumulation
        r2t[b, : int(lens[b])] = s.to(torch.int32)
        means.append(s)
    v_vals = (torch.arange(T) % 128).float()  # slot s -> value s % 128 (exact in bf16)
    k_buf = torch.zeros(T, Hkv, D, dtype=torch.bfloat16, device="cuda")
    v_buf = v_vals.view(T, 1, 1).expand(T, Hkv, D).contiguous().to(torch.bfloat16).cuda()
    q = torch.randn(B, Hq, D, generator=g).to(torch.bfloat16).cuda()
    out = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device="cuda")
    ops.decode_attention(q, k_buf, v_buf, out, None, None, None, None, r2t.cuda(), torch.arange(B).cuda(), lens.cuda(),
                         0.1, 0.0, max_seq_len=256)
    for b in range(B):
        exp = (means[b] % 128).double().mean()
        got = out[b].float().cpu()
        assert (got - exp).abs().max() <= abs(exp) * 2 ** -8 + 1e-6


@pytest.mark.parametrize("cfg", ["cfg1", "cfg2_slice", "cfg4_tp8_slice"])
def test_baseline_config_shapes(cfg):
    """BASELINE.json configs at (or close to) full size, compared with the oracle on a few requests, plus
    size-independent properties on the whole batch."""
    from iaas_sglang_b200 import ops

    if cfg == "cfg1":
        check(DecodeCase(16, 32, 8, 64, 64, 1024), fused=True)
        return
    if cfg == "cfg2_slice":
        B, Hq, Hkv, D, S, shared = 64, 32, 8, 128, 4096, 0
    else:
        B, Hq, Hkv, D, S, shared = 256, 8, 1, 128, 8192, 2048
    c = DecodeCase(B, Hq, Hkv, D, D, S, shared_prefix=shared)
    d = c.cuda()
    out = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device="cuda")
    args = (d["k_new"], d["v_new"], d["loc"], None, d["req_to_token"], d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0)
    ops.decode_attention(d["q"], d["k_buf"], d["v_buf"], out, *args, max_seq_len=S, fused_kv_write=True)
    # (a) oracle on a sample of requests (the oracle reads only the rows those requests reference)
    sample = [0, B // 2, B - 1]
    sub = DecodeCase.__new__(DecodeCase)
    sub.__dict__.update(c.__dict__)
    sub.B = len(sample)
    sub.q, sub.k_new, sub.v_new = c.q[sample], c.k_new[sample], c.v_new[sample]
    sub.req_pool_indices, sub.seq_lens, sub.loc = c.req_pool_indices[sample], c.seq_lens[sample], c.loc[sample]
    ref, _, _ = sub.oracle()
    assert rel_l2(out[sample].cpu(), ref) < TOL
    # (b) split invariance: 1 split vs the heuristic
    out1 = torch.zeros_like(out)
    ops.decode_attention(d["q"], d["k_buf"], d["v_buf"], out1, *args, max_seq_len=S, num_splits=1)
    assert rel_l2(out1, out) < 1e-3
    # (c) linearity in V with a power-of-two factor is bit-exact
    out2 = torch.zeros_like(out)
    ops.decode_attention(d["q"], d["k_buf"], d["v_buf"] * 2, out2, d["k_new"], d["v_new"] * 2, *args[2:], max_seq_len=S)
    assert torch.equal(out2, out * 2)
    # (d) appended rows landed exactly
    assert torch.equal(d["k_buf"][d["loc"]], d["k_new"]) and torch.equal(d["v_buf"][d["loc"]], d["v_new"])


# ---- MLA absorbed decode, tuned kernel (decode_mla.cuh): 16-head q blocks over 64-token tiles ----
@pytest.mark.parametrize("seq_lens", [[1], [1, 2, 3, 17], [63, 64, 65, 31, 32, 33], [128, 129, 127, 1], [5000, 3, 977],
                                      [2048] * 5])
@pytest.mark.parametrize("Hq", [16, 22, 5, 128])
def test_mla_ragged_lengths_and_head_counts(seq_lens, Hq):
    check(DecodeCase(len(seq_lens), Hq, 1, 576, 512, seq_lens), fused=True)
    check(DecodeCase(len(seq_lens), Hq, 1, 576, 512, seq_lens), fused=False)


def test_mla_fp16_no_write_and_padding_rows():
    c = DecodeCase(3, 16, 1, 576, 512, [700, 64, 9], dtype=torch.float16)
    ref, _, _ = c.oracle(round_p=False)
    out, _, _ = run_cuda_decode(c, fused=True)
    assert rel_l2(out, ref) < 5e-4
    check(DecodeCase(2, 16, 1, 576, 512, 300), write_kv=False)
    c = DecodeCase(4, 16, 1, 576, 512, [300, 1, 1, 1])
    c.req_pool_indices[1:] = 0
    c.loc[1:] = 0
    check(c, first_slot=1, fused=True)


def test_mla_tuned_kernel_matches_generic_kernel_and_many_ctas_per_request():
    """One long request is split over all 148 CTAs (partials + merge); the generic kernel (explicit split count) is
    an independent implementation of the same op."""
    c = DecodeCase(2, 16, 1, 576, 512, [40000, 150])
    a, _, _ = run_cuda_decode(c, fused=True)
    b, _, _ = run_cuda_decode(c, fused=True, num_splits=16)
    assert rel_l2(a, b) < 1e-3


def test_mla_cfg5_tp8_slice_properties():
    """BASELINE config 5 per-GPU shape (16 heads, latent 576/512) on a slice of the batch: oracle on a sample,
    exact linearity in the latent's value part, exact append."""
    from iaas_sglang_b200 import ops

    B, Hq, S = 24, 16, 16384
    c = DecodeCase(B, Hq, 1, 576, 512, S)
    d = c.cuda()
    out = torch.zeros(B, Hq, 512, dtype=torch.bfloat16, device="cuda")
    tail = (d["loc"], None, d["req_to_token"], d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0)
    ops.decode_attention(d["q"], d["k_buf"], d["v_buf"], out, d["k_new"], d["v_new"], *tail, max_seq_len=S,
                         fused_kv_write=True)
    sample = [0, B // 2, B - 1]
    sub = DecodeCase.__new__(DecodeCase)
    sub.__dict__.update(c.__dict__)
    sub.B = len(sample)
    sub.q, sub.k_new, sub.v_new = c.q[sample], c.k_new[sample], c.v_new[sample]
    sub.req_pool_indices, sub.seq_lens, sub.loc = c.req_pool_indices[sample], c.seq_lens[sample], c.loc[sample]
    ref, _, _ = sub.oracle(round_p=False)
    assert rel_l2(out[sample].cpu(), ref) < TOL
    assert torch.equal(d["k_buf"][d["loc"]], d["k_new"])
    # scaling the value columns (first 512) by 2 would also change the keys, so scale a copy used as V only: the
    # generic kernel accepts distinct K and V pools and must agree with the tuned one
    out_g = torch.zeros_like(out)
    ops.decode_attention(d["q"], d["k_buf"], d["v_buf"], out_g, None, None, None, None, *tail[2:], max_seq_len=S,
                         num_splits=8)
    assert rel_l2(out_g, out) < 1e-3


@pytest.mark.parametrize("seq_lens", [[4096] * 8, [1, 2, 3, 17], [5000, 3, 977, 64, 65], [300] * 64, [0, 7, 4096]])
@pytest.mark.parametrize("heads", [(32, 8, 128), (8, 1, 128), (20, 4, 128), (32, 8, 64)],
                         ids=lambda h: f"Hq{h[0]}Hkv{h[1]}D{h[2]}")
def test_fused_fp8_quant_of_the_output_is_bit_identical(seq_lens, heads):
    """quant_fp8=True: (x_q, x_s) from the split-merge kernel (Dv == 128; two launches otherwise) must equal
    per_token_group_quant_fp8(out.view(B, Hq*Dv), 128, column_major_scales=True) bit for bit, and `out` itself must
    not change (fp8_utils.py:202-204 is the call this replaces)."""
    from iaas_sglang_b200 import ops

    Hq, Hkv, D = heads
    lens = seq_lens
    c = DecodeCase(len(lens), Hq, Hkv, D, D, [max(s, 1) for s in lens])
    d = c.cuda("cuda")
    if 0 in lens:  # padded CUDA-graph rows have length 0: defined zeros, scale eps / 448
        d["seq_lens"] = torch.tensor(lens, dtype=torch.int64, device="cuda")
    kb0, vb0 = d["k_buf"].clone(), d["v_buf"].clone()
    args = lambda kb, vb, out: (d["q"], kb, vb, out, d["k_new"], d["v_new"], d["loc"], None, d["req_to_token"],  # noqa: E731
                                d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0)
    out_a = torch.zeros(c.B, Hq, D, dtype=c.dtype, device="cuda")
    ops.decode_attention(*args(d["k_buf"], d["v_buf"], out_a), max_seq_len=max(lens), fused_kv_write=True)
    q_ref, s_ref = ops.per_token_group_quant_fp8(out_a.view(c.B, Hq * D), 128, column_major_scales=True)
    out_b = torch.full((c.B, Hq, D), 7.0, dtype=c.dtype, device="cuda")
    res = ops.decode_attention(*args(kb0, vb0, out_b), max_seq_len=max(lens), fused_kv_write=True, quant_fp8=True)
    torch.cuda.synchronize()
    assert res[0] is out_b
    assert torch.equal(out_a, out_b)
    assert torch.equal(res[1].view(torch.uint8), q_ref.view(torch.uint8))
    assert torch.equal(res[2], s_ref) and res[2].stride() == s_ref.stride()
