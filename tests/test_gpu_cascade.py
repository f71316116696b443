"""GPU tier: cascade (shared-prefix) decode -- merge_state, the shared-prefix detector and the cascade path of the decode
attention op against the CPU oracle.  The cascade path must give the SAME result as plain decode attention (the oracle
knows nothing about prefixes): it only reads the shared KV rows once."""
import pytest
import torch

from oracle import cpu_oracle as orc
from tests.util import DecodeCase, rel_l2

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
def test_merge_state_against_oracle(dtype):
    from iaas_sglang_b200 import ops

    g = torch.Generator().manual_seed(3)
    T, H, D = 37, 12, 128
    va, vb = torch.randn(T, H, D, generator=g).to(dtype), torch.randn(T, H, D, generator=g).to(dtype)
    sa, sb = torch.randn(T, H, generator=g) * 5, torch.randn(T, H, generator=g) * 5
    sa[0, 0], sb[1, 1] = float("-inf"), float("inf")      # empty states (merge_attn_states.cu:66-67)
    sa[2, 2], sb[2, 2] = float("-inf"), float("-inf")     # both empty
    ref_v, ref_s = orc.merge_state(va, sa, vb, sb)
    out_v, out_s = ops.merge_state(va.cuda(), sa.cuda(), vb.cuda(), sb.cuda())
    assert rel_l2(out_v, ref_v) < 2e-3
    ok = ~torch.isinf(ref_s)
    torch.testing.assert_close(out_s.cpu()[ok], ref_s[ok], atol=1e-5, rtol=1e-5)
    assert torch.equal(torch.isinf(out_s.cpu()), torch.isinf(ref_s))
    v2, s2 = ops.merge_state_v2(va.cuda(), sa.cuda(), vb.cuda(), sb.cuda())   # same contract (attention.py:31-52)
    assert torch.equal(v2, out_v) and torch.equal(s2.isinf(), out_s.isinf())


@pytest.mark.parametrize("shared,lens,expect", [(300, [900, 400, 301, 1000], 256), (128, [500, 600], 0), (512, [513, 2000, 700], 512),
                                                (0, [400, 400], 0), (700, [1200], 0), (1024, [1025] * 5, 1024)])
@pytest.mark.parametrize("meta", [torch.int64, torch.int32], ids=["i64", "i32"])
def test_shared_prefix_len_bit_exact(shared, lens, expect, meta):
    from iaas_sglang_b200 import ops

    c = DecodeCase(len(lens), 8, 1, 128, 128, lens, shared_prefix=shared, extra_reqs=2, seed=5)
    want = orc.shared_prefix_len(c.req_to_token, c.req_pool_indices, c.seq_lens)
    assert want == expect
    st = ops.CascadeState(len(lens), 8, 128, torch.bfloat16, "cuda")
    for _ in range(2):   # the scratch re-arms itself
        st.detect(c.req_to_token.cuda(), c.req_pool_indices.to(meta).cuda(), c.seq_lens.to(meta).cuda())
        assert int(st.prefix_len.item()) == want
        assert int(st.one_seq_len.item()) == want + len(lens)
        assert int(st.one_req.item()) == int(c.req_pool_indices[0])


CASES = [
    # B, Hq, Hkv, D, lens, shared prefix
    (4, 8, 1, 128, [2500, 2100, 4000, 2049], 2048),      # cfg 4's TP8-local shape, small
    (6, 32, 8, 128, [700, 300, 513, 1024, 290, 2000], 280),   # prefix not a multiple of the alignment: 256 go the cascade way
    (3, 16, 2, 64, [1500, 1301, 1400], 1300),            # (a request's NEW token never sits inside the shared prefix: allocator contract)
    (5, 8, 8, 128, [600, 700, 800, 650, 900], 0),        # nothing shared: P = 0, the prefix state is empty
    (1, 8, 2, 128, [3000], 0),                           # a batch of one
]


@pytest.mark.parametrize("B,Hq,Hkv,D,lens,shared", CASES)
@pytest.mark.parametrize("fused", [False, True], ids=["sep-append", "fused-append"])
def test_cascade_decode_equals_plain_decode(B, Hq, Hkv, D, lens, shared, fused):
    from iaas_sglang_b200 import ops

    c = DecodeCase(B, Hq, Hkv, D, D, lens, shared_prefix=shared, extra_reqs=1, seed=B + Hq)
    ref, kb, vb = c.oracle()
    d = c.cuda()
    st = ops.CascadeState(B, Hq, D, torch.bfloat16, "cuda")
    st.detect(d["req_to_token"], d["req_pool_indices"], d["seq_lens"])
    assert int(st.prefix_len.item()) == orc.shared_prefix_len(c.req_to_token, c.req_pool_indices, c.seq_lens)
    st.prefix_attention(d["q"], d["k_buf"], d["v_buf"], d["req_to_token"], c.sm_scale, 0.0)
    out = torch.empty(B, Hq, D, dtype=torch.bfloat16, device="cuda")
    ops.decode_attention(d["q"], d["k_buf"], d["v_buf"], out, d["k_new"], d["v_new"], d["loc"], None, d["req_to_token"],
                         d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0, fused_kv_write=fused, cascade=st)
    assert rel_l2(out, ref) < 2e-3, rel_l2(out, ref)
    assert torch.equal(d["k_buf"].cpu()[1:], kb[1:]) and torch.equal(d["v_buf"].cpu()[1:], vb[1:])   # the append still happens
    # and the fused FP8 epilogue sees the merged rows
    if D == 128:
        o2 = torch.empty_like(out)
        st.prefix_attention(d["q"], d["k_buf"], d["v_buf"], d["req_to_token"], c.sm_scale, 0.0)
        _, xq, xs = ops.decode_attention(d["q"], d["k_buf"], d["v_buf"], o2, None, None, None, None, d["req_to_token"],
                                         d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0, quant_fp8=True, cascade=st)
        q0, s0 = ops.per_token_group_quant_fp8(o2.view(B, Hq * D), 128, column_major_scales=True)
        assert torch.equal(xq.view(torch.uint8), q0.view(torch.uint8)) and torch.equal(xs, s0)
        assert rel_l2(o2, ref) < 2e-3


def test_plugin_cascade_eager_and_cuda_graph():
    """B200AttnBackend(cascade=True): init_forward_metadata finds the shared prefix on the device, forward_decode runs the
    prefix kernel + suffix kernel + merge; a CUDA graph captured on dummy metadata (no prefix) follows the real batch at
    replay (the prefix length lives in device memory)."""
    from iaas_sglang_b200 import compat
    from iaas_sglang_b200.backend import B200AttnBackend

    B, Hq, Hkv, D, ctx = 4, 16, 2, 128, 3000
    lens = [3000, 1400, 2222, 1300]
    c = DecodeCase(B, Hq, Hkv, D, D, lens, max_ctx=ctx, shared_prefix=1290, extra_reqs=3, seed=21)
    ref, kb, vb = c.oracle()
    dev = torch.device("cuda", 0)
    T = c.k_buf.shape[0]

    def fresh():
        r2t = compat.ReqToTokenPool(B + 3, ctx, dev)
        kv = compat.MHATokenToKVPool(T, 1, torch.bfloat16, Hkv, D, 1, dev)
        runner = compat.ModelRunnerStub(compat.ModelConfigStub(Hq, Hkv, D, ctx), r2t, kv, device=dev)
        be = B200AttnBackend(runner, cascade=True)
        kv.k_buffer[0][:T].copy_(c.k_buf)
        kv.v_buffer[0][:T].copy_(c.v_buf)
        return r2t, kv, be

    layer = compat.RadixAttention(Hq, D, c.sm_scale, Hkv, 0)
    # ---- eager ----
    r2t, kv, be = fresh()
    r2t.req_to_token.copy_(c.req_to_token)
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, B, req_pool_indices=c.req_pool_indices.cuda(), seq_lens=c.seq_lens.cuda(),
                             out_cache_loc=c.loc.cuda(), seq_lens_sum=int(c.seq_lens.sum()), seq_lens_cpu=c.seq_lens,
                             req_to_token_pool=r2t, token_to_kv_pool=kv, attn_backend=be)
    be.init_forward_metadata(fb)
    assert int(be.cascade_state.prefix_len.item()) == 1280
    o = layer(c.q.cuda().view(B, -1), c.k_new.cuda().view(B, -1), c.v_new.cuda().view(B, -1), fb)
    assert rel_l2(o.view(B, Hq, D), ref) < 2e-3
    assert torch.equal(kv.k_buffer[0][:T].cpu()[1:], kb[1:])
    # ---- CUDA graph: capture on dummies, replay on the real batch ----
    r2t, kv, be = fresh()
    be.init_cuda_graph_state(max_bs=B, max_num_tokens=B)
    seq_lens = torch.full((B,), be.get_cuda_graph_seq_len_fill_value(), dtype=torch.int32, device=dev)
    rpi = torch.zeros(B, dtype=torch.int32, device=dev)
    loc = torch.zeros(B, dtype=torch.int64, device=dev)
    q = torch.zeros(B, Hq * D, dtype=torch.bfloat16, device=dev)
    k = torch.zeros(B, Hkv * D, dtype=torch.bfloat16, device=dev)
    v = torch.zeros(B, Hkv * D, dtype=torch.bfloat16, device=dev)
    out = torch.zeros(B, Hq * D, dtype=torch.bfloat16, device=dev)
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, B, req_pool_indices=rpi, seq_lens=seq_lens, out_cache_loc=loc,
                             req_to_token_pool=r2t, token_to_kv_pool=kv, attn_backend=be)
    be.init_forward_metadata_capture_cuda_graph(B, B, rpi, seq_lens, None, compat.ForwardMode.DECODE, None)
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        layer(q, k, v, fb)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            out.copy_(layer(q, k, v, fb))
    r2t.req_to_token.copy_(c.req_to_token)
    seq_lens.copy_(c.seq_lens); rpi.copy_(c.req_pool_indices); loc.copy_(c.loc)
    q.copy_(c.q.view(B, -1)); k.copy_(c.k_new.view(B, -1)); v.copy_(c.v_new.view(B, -1))
    be.init_forward_metadata_replay_cuda_graph(B, rpi, seq_lens, int(c.seq_lens.sum()), None, compat.ForwardMode.DECODE, None,
                                               seq_lens_cpu=c.seq_lens)
    graph.replay()
    torch.cuda.synchronize()
    assert int(be.cascade_state.prefix_len.item()) == 1280
    assert rel_l2(out.view(B, Hq, D), ref) < 2e-3
