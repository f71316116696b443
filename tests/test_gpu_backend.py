"""GPU tier: the AttentionBackend plugin surface (B200AttnBackend) driven the way ModelRunner drives a backend
(python/sglang/test/attention/test_flashattn_backend.py:15-140 is the harness pattern), extend attention, the
page-table index builders, CUDA-graph capture/replay and the steps either side of the path."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as orc
from tests.util import DecodeCase, golden, golden_extend_cases, rel_l2

pytestmark = pytest.mark.gpu


def make_runner(B, Hq, Hkv, D, ctx, layers=2, mla=False, dtype=torch.bfloat16):
    from iaas_sglang_b200 import compat
    from iaas_sglang_b200.backend import B200AttnBackend

    dev = torch.device("cuda", 0)
    T = B * ctx
    r2t = compat.ReqToTokenPool(B + 3, ctx, dev)
    if mla:
        kv = compat.MLATokenToKVPool(T, 1, dtype, 512, 64, layers, dev)
        mc = compat.ModelConfigStub(Hq, 1, 576, ctx, v_head_dim=512, attention_arch_is_mla=True, kv_lora_rank=512, qk_rope_head_dim=64)
    else:
        kv = compat.MHATokenToKVPool(T, 1, dtype, Hkv, D, layers, dev)
        mc = compat.ModelConfigStub(Hq, Hkv, D, ctx)
    runner = compat.ModelRunnerStub(mc, r2t, kv, device=dev)
    return runner, B200AttnBackend(runner)


def test_forward_decode_through_plugin_matches_oracle():
    from iaas_sglang_b200 import compat

    B, Hq, Hkv, D, ctx = 5, 32, 8, 128, 700
    runner, backend = make_runner(B, Hq, Hkv, D, ctx)
    c = DecodeCase(B, Hq, Hkv, D, D, [700, 1, 64, 333, 512], max_ctx=ctx, extra_reqs=3)
    pool, r2t = runner.token_to_kv_pool, runner.req_to_token_pool
    r2t.req_to_token.copy_(c.req_to_token)
    T = c.k_buf.shape[0]
    for l in range(2):
        pool.k_buffer[l][:T].copy_(c.k_buf)
        pool.v_buffer[l][:T].copy_(c.v_buf)
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, B, req_pool_indices=c.req_pool_indices.cuda(), seq_lens=c.seq_lens.cuda(),
                             out_cache_loc=c.loc.cuda(), seq_lens_sum=int(c.seq_lens.sum()), seq_lens_cpu=c.seq_lens,
                             req_to_token_pool=r2t, token_to_kv_pool=pool, attn_backend=backend)
    backend.init_forward_metadata(fb)
    ref, kb, vb = c.oracle()
    for l in range(2):
        layer = compat.RadixAttention(Hq, D, c.sm_scale, Hkv, l)
        o = layer(c.q.cuda().view(B, Hq * D), c.k_new.cuda().view(B, Hkv * D), c.v_new.cuda().view(B, Hkv * D), fb)
        assert o.shape == (B, Hq * D) and o.dtype == torch.bfloat16
        assert rel_l2(o.view(B, Hq, D), ref) < 2e-3
        assert torch.equal(pool.k_buffer[l][:T].cpu(), kb) and torch.equal(pool.v_buffer[l][:T].cpu(), vb)
    # save_kv_cache=False must not touch the pool
    before = pool.k_buffer[0].clone()
    layer = compat.RadixAttention(Hq, D, c.sm_scale, Hkv, 0)
    backend.forward_decode(c.q.cuda(), torch.zeros_like(c.k_new).cuda(), torch.zeros_like(c.v_new).cuda(), layer, fb, save_kv_cache=False)
    assert torch.equal(before, pool.k_buffer[0])


def test_mla_forward_decode_through_plugin():
    from iaas_sglang_b200 import compat

    B, Hq, ctx = 3, 16, 400
    runner, backend = make_runner(B, Hq, 1, 576, ctx, layers=1, mla=True)
    c = DecodeCase(B, Hq, 1, 576, 512, [400, 33, 256], max_ctx=ctx, extra_reqs=3)
    pool, r2t = runner.token_to_kv_pool, runner.req_to_token_pool
    r2t.req_to_token.copy_(c.req_to_token)
    T = c.k_buf.shape[0]
    pool.kv_buffer[0][:T].copy_(c.k_buf)
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, B, req_pool_indices=c.req_pool_indices.cuda(), seq_lens=c.seq_lens.cuda(),
                             out_cache_loc=c.loc.cuda(), seq_lens_sum=int(c.seq_lens.sum()), seq_lens_cpu=c.seq_lens,
                             req_to_token_pool=r2t, token_to_kv_pool=pool, attn_backend=backend)
    backend.init_forward_metadata(fb)
    layer = compat.RadixAttention(Hq, 576, c.sm_scale, 1, 0, v_head_dim=512)
    o = backend.forward_decode(c.q.cuda(), c.k_new.cuda(), c.v_new.cuda(), layer, fb)
    ref, kb, _ = c.oracle(round_p=False)  # exact softmax; see TOL_MLA_LITERAL in test_gpu_decode_attention.py
    assert o.shape == (B, Hq * 512)
    assert rel_l2(o.view(B, Hq, 512), ref) < 2e-3
    assert torch.equal(pool.kv_buffer[0][:T].cpu(), kb)


def test_cuda_graph_capture_and_replay():
    """cuda_graph_runner.py:456-703 contract: capture with padded/dummy metadata in fixed buffers, then replay after
    only the contents of seq_lens / req_pool_indices / out_cache_loc / req_to_token changed."""
    from iaas_sglang_b200 import compat

    B, Hq, Hkv, D, ctx = 4, 32, 8, 128, 2048
    runner, backend = make_runner(B, Hq, Hkv, D, ctx, layers=1)
    runner.model_config.context_len = ctx
    backend.init_cuda_graph_state(max_bs=B, max_num_tokens=B)
    c = DecodeCase(B, Hq, Hkv, D, D, [2048, 5, 777, 1500], max_ctx=ctx, extra_reqs=3)
    pool, r2t = runner.token_to_kv_pool, runner.req_to_token_pool
    T = c.k_buf.shape[0]
    pool.k_buffer[0][:T].copy_(c.k_buf)
    pool.v_buffer[0][:T].copy_(c.v_buf)
    seq_lens = torch.full((B,), backend.get_cuda_graph_seq_len_fill_value(), dtype=torch.int64, device="cuda")
    rpi = torch.zeros(B, dtype=torch.int64, device="cuda")
    loc = torch.zeros(B, dtype=torch.int64, device="cuda")
    q = torch.zeros(B, Hq * D, dtype=torch.bfloat16, device="cuda")
    k = torch.zeros(B, Hkv * D, dtype=torch.bfloat16, device="cuda")
    v = torch.zeros(B, Hkv * D, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(B, Hq * D, dtype=torch.bfloat16, device="cuda")
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, B, req_pool_indices=rpi, seq_lens=seq_lens, out_cache_loc=loc,
                             req_to_token_pool=r2t, token_to_kv_pool=pool, attn_backend=backend)
    layer = compat.RadixAttention(Hq, D, c.sm_scale, Hkv, 0)
    backend.init_forward_metadata_capture_cuda_graph(B, B, rpi, seq_lens, None, compat.ForwardMode.DECODE, None)
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        layer(q, k, v, fb)  # warm-up outside the graph (lazy cudaFuncSetAttribute)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            out.copy_(layer(q, k, v, fb))
    # new step: only buffer CONTENTS change
    r2t.req_to_token.copy_(c.req_to_token)
    seq_lens.copy_(c.seq_lens)
    rpi.copy_(c.req_pool_indices)
    loc.copy_(c.loc)
    q.copy_(c.q.view(B, -1))
    k.copy_(c.k_new.view(B, -1))
    v.copy_(c.v_new.view(B, -1))
    backend.init_forward_metadata_replay_cuda_graph(B, rpi, seq_lens, int(c.seq_lens.sum()), None, compat.ForwardMode.DECODE, None,
                                                    seq_lens_cpu=c.seq_lens)
    graph.replay()
    torch.cuda.synchronize()
    ref, kb, vb = c.oracle()
    assert rel_l2(out.view(B, Hq, D), ref) < 2e-3
    assert torch.equal(pool.k_buffer[0][:T].cpu()[1:], kb[1:])  # slot 0 took the capture-time dummy write


def test_unsupported_modes_raise():
    from iaas_sglang_b200 import compat
    from iaas_sglang_b200.backend import B200AttnBackend

    runner, backend = make_runner(2, 8, 2, 64, 64, layers=1)
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, 2, spec_info=object())
    with pytest.raises(NotImplementedError):
        backend.init_forward_metadata(fb)
    runner.sliding_window_size = 128
    with pytest.raises(NotImplementedError):
        B200AttnBackend(runner)


@pytest.mark.parametrize("case", list(golden_extend_cases()), ids=lambda c: f"c{c['idx']}")
def test_extend_attention_against_reference_golden(case):
    from iaas_sglang_b200 import ops

    c = case
    d = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in c.items()}
    out = torch.zeros_like(d["o"])
    ops.extend_attention(d["q"], d["k_ext"], d["v_ext"], out, d["k_buf"], d["v_buf"], d["r2t"], d["rpi"], d["seq_lens"],
                         d["ext_lens"], d["ext_start"], int(c["ext_lens"].max()), 1.0 / c["D"] ** 0.5, 0.0)
    torch.testing.assert_close(out.float().cpu(), c["o"].float(), atol=1e-2, rtol=1e-2)  # test_extend.py:183
    assert rel_l2(out, c["o"]) < 5e-3


def test_forward_extend_through_plugin_then_decode():
    """prefill 3 requests through forward_extend (writes KV), then one decode step over the same pool."""
    from iaas_sglang_b200 import compat

    B, Hq, Hkv, D, ctx = 3, 16, 4, 128, 96
    runner, backend = make_runner(B, Hq, Hkv, D, ctx, layers=1)
    pool, r2t = runner.token_to_kv_pool, runner.req_to_token_pool
    g = torch.Generator().manual_seed(5)
    ext = torch.tensor([40, 7, 64], dtype=torch.int64)
    T = int(ext.sum())
    slots = torch.randperm(B * ctx - 1, generator=g)[:T] + 1
    start = torch.zeros(B, dtype=torch.int64)
    start[1:] = torch.cumsum(ext[:-1], 0)
    table = torch.zeros(B + 3, ctx, dtype=torch.int32)
    for b in range(B):
        table[b, : ext[b]] = slots[start[b]: start[b] + ext[b]].to(torch.int32)
    r2t.req_to_token.copy_(table)
    q = torch.randn(T, Hq * D, generator=g).to(torch.bfloat16)
    k = torch.randn(T, Hkv * D, generator=g).to(torch.bfloat16)
    v = torch.randn(T, Hkv * D, generator=g).to(torch.bfloat16)
    fb = compat.ForwardBatch(compat.ForwardMode.EXTEND, B, req_pool_indices=torch.arange(B).cuda(), seq_lens=ext.cuda(),
                             out_cache_loc=slots.cuda(), extend_seq_lens=ext.cuda(), extend_start_loc=start.cuda(),
                             extend_prefix_lens=torch.zeros(B, dtype=torch.int64).cuda(), extend_seq_lens_cpu=ext.tolist(),
                             req_to_token_pool=r2t, token_to_kv_pool=pool, attn_backend=backend)
    backend.init_forward_metadata(fb)
    layer = compat.RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    o = layer(q.cuda(), k.cuda(), v.cuda(), fb)
    kb = torch.zeros(B * ctx + 1, Hkv, D, dtype=torch.bfloat16)
    vb = torch.zeros_like(kb)
    kb[slots], vb[slots] = k.view(T, Hkv, D), v.view(T, Hkv, D)
    assert torch.equal(pool.k_buffer[0].cpu(), kb) and torch.equal(pool.v_buffer[0].cpu(), vb)
    ref = torch.zeros(T, Hq, D, dtype=torch.bfloat16)
    orc.extend_attention(q.view(T, Hq, D), k.view(T, Hkv, D), v.view(T, Hkv, D), ref, kb, vb, table, torch.arange(B), ext, ext,
                         start, int(ext.max()), D ** -0.5, 0.0)
    assert rel_l2(o.view(T, Hq, D), ref) < 5e-3


def test_kv_index_builders_bit_exact():
    from iaas_sglang_b200 import ops

    g = golden("kv_indices")
    r2t = torch.from_numpy(g["r2t"]).cuda()
    P = int(g["page_size"])
    for i in range(int(g["n_cases"])):
        p = f"c{i}_"
        rpi, lens, indptr = (torch.from_numpy(g[p + n]).cuda() for n in ("rpi", "lens", "indptr"))
        out = torch.full((int(indptr[-1]),), -7, dtype=torch.int32, device="cuda")
        ops.create_kv_indices(r2t, rpi, lens, indptr, None, out)
        assert np.array_equal(out.cpu().numpy(), g[p + "kv_indices"])
        out2 = torch.full((int(g[p + "indptr2"][-1]),), -7, dtype=torch.int32, device="cuda")
        ops.create_kv_indices(r2t, rpi, torch.from_numpy(g[p + "lens2"]).cuda(), torch.from_numpy(g[p + "indptr2"]).cuda(),
                              torch.from_numpy(g[p + "start"]).cuda(), out2)
        assert np.array_equal(out2.cpu().numpy(), g[p + "kv_indices2"])
        pt = torch.full(g[p + "page_table"].shape, -1, dtype=torch.int32, device="cuda")
        ops.create_page_table(r2t, rpi, lens, None, pt, P)
        assert np.array_equal(pt.cpu().numpy(), g[p + "page_table"])
    # test_create_kvindices.py:66-71 sizes (batch 1786 of 4096 x 4096), against the reference expression
    rng = np.random.RandomState(0)
    big = torch.arange(4096 * 4096, dtype=torch.int32, device="cuda").view(4096, 4096)
    rpi = torch.from_numpy(rng.choice(4096, 1786, replace=False)).cuda()
    lens = torch.from_numpy(rng.choice(4096, 1786, replace=False)).cuda()
    indptr = torch.zeros(1787, dtype=torch.int64, device="cuda")
    indptr[1:] = torch.cumsum(lens, 0)
    out = torch.empty(int(indptr[-1]), dtype=torch.int32, device="cuda")
    ops.create_kv_indices(big, rpi, lens, indptr, None, out)
    ref = torch.cat([big[rpi[j], : lens[j]] for j in range(1786)])
    assert torch.equal(out, ref)


def test_set_kv_buffer_bit_exact():
    from iaas_sglang_b200 import ops

    kb = torch.zeros(100, 8, 128, dtype=torch.bfloat16, device="cuda")
    vb = torch.zeros(100, 8, 128, dtype=torch.bfloat16, device="cuda")
    k, v = torch.randn(7, 8, 128).bfloat16().cuda(), torch.randn(7, 8, 128).bfloat16().cuda()
    loc = torch.tensor([5, 99, 0, 17, 3, 42, 1]).cuda()
    ops.set_kv_buffer(kb, vb, loc, k, v)
    ek, ev = torch.zeros_like(kb), torch.zeros_like(vb)
    ek[loc], ev[loc] = k, v
    assert torch.equal(kb, ek) and torch.equal(vb, ev)


def test_steps_either_side_of_the_path():
    from iaas_sglang_b200 import ops

    x = torch.randn(64, 4096).bfloat16()
    w = (torch.rand(4096) + 0.5).bfloat16()
    assert rel_l2(ops.rmsnorm(x.cuda(), w.cuda(), 1e-5), orc.rmsnorm(x, w, 1e-5)) < 4e-3
    res = torch.randn(64, 4096).bfloat16()
    res_d = res.cuda()
    out = ops.rmsnorm(x.cuda(), w.cuda(), 1e-5, residual=res_d)
    summed = (x.float() + res.float()).bfloat16()
    assert torch.equal(res_d.cpu(), summed)
    assert rel_l2(out, orc.rmsnorm(summed, w, 1e-5)) < 4e-3
    y = torch.randn(64, 2 * 14336).bfloat16()
    assert rel_l2(ops.silu_and_mul(y.cuda()), orc.silu_and_mul(y)) < 4e-3
    # neox rope against the textbook form (rotary_embedding.py _apply_rotary_emb, neox style)
    M, Hq, Hkv, D = 9, 32, 8, 128
    q, k = torch.randn(M, Hq, D).bfloat16(), torch.randn(M, Hkv, D).bfloat16()
    pos = torch.randint(0, 8192, (M,), dtype=torch.int64)
    inv = 1.0 / (500000.0 ** (torch.arange(0, D, 2).float() / D))
    ang = pos.float()[:, None] * inv[None]
    cos, sin = ang.cos()[:, None, :], ang.sin()[:, None, :]

    def rot(t):
        t1, t2 = t.float()[..., : D // 2], t.float()[..., D // 2:]
        return torch.cat([t1 * cos - t2 * sin, t2 * cos + t1 * sin], -1)

    qd, kd = q.cuda(), k.cuda()
    ops.rope_neox_(qd, kd, pos.cuda(), D, 500000.0)
    assert rel_l2(qd, rot(q)) < 4e-3 and rel_l2(kd, rot(k)) < 4e-3
