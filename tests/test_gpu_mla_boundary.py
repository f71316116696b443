"""GPU tier: the MLA boundary of SURVEY 8a row a7 -- ops.cutlass_mla_decode with the reference's signature
(sgl-kernel/python/sgl_kernel/attention.py:54-138) on the inputs of the reference's own test
(sgl-kernel/tests/test_cutlass_mla.py:38-99), the q_rope= / k_rope= arguments of
cutlass_mla_backend.py:226-285 through the plugin, and the int32 metadata of cuda_graph_runner.py:273-276.

Tolerances: the reference test's own bar (atol = rtol = 1e-2, test_cutlass_mla.py:99) against the literal SDPA
reference, and 2e-3 relative L2 (BASELINE.json north_star) against the CPU oracle with the exact softmax."""
import pytest
import torch

from oracle import cpu_oracle as orc
from tests.util import DecodeCase, rel_l2

pytestmark = pytest.mark.gpu


def ref_mla(query, kv_cache, scale, block_tables, seq_lens, dv=512):
    """ref_mla of test_cutlass_mla.py:14-35 in fp32 (gather the pages, softmax(q k^T) v per request)."""
    bs, H, D = query.shape
    out = torch.zeros(bs, H, dv, dtype=torch.float32, device=query.device)
    for i in range(bs):
        kv = kv_cache[block_tables[i].long()].view(-1, D)[: int(seq_lens[i])].float()
        s = (query[i].float() @ kv.t()) * scale
        out[i] = torch.softmax(s, dim=-1) @ kv[:, :dv]
    return out


def page_table_to_token_slots(page_table, P, max_len):
    """token-granular req_to_token equivalent of a page table (memory_pool.py: slot = page * page_size + offset)."""
    t = torch.arange(max_len)
    return (page_table[:, t // P].long() * P + (t % P)[None, :]).to(torch.int32)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
@pytest.mark.parametrize("mean_seq_len", [128, 1024, 4096])
@pytest.mark.parametrize("bs,varlen", [(1, False), (2, True), (4, True), (4, False)])
@pytest.mark.parametrize("block_size", [1, 16, 64, 128])
@pytest.mark.parametrize("num_heads", [16, 32, 128])
def test_cutlass_mla_decode_on_the_reference_test_inputs(dtype, mean_seq_len, bs, varlen, block_size, num_heads):
    from iaas_sglang_b200 import ops

    torch.manual_seed(42)
    d, dv, dev = 576, 512, "cuda"
    scale = (128 + 64) ** (-0.5)
    if varlen:
        seq_lens = torch.empty(bs).normal_(mean_seq_len, mean_seq_len / 2).clip(2).to(torch.int32)
    else:
        seq_lens = torch.full((bs,), mean_seq_len, dtype=torch.int32)
    max_seq_len = int(seq_lens.max())
    pack = 128 // block_size
    block_num = ((max_seq_len + block_size - 1) // block_size + pack - 1) // pack * pack
    q = (torch.randn(bs, num_heads, d) * 100.0).to(dtype).to(dev)     # "larger q values to detect split kv error"
    block_table = torch.randint(0, bs * block_num, (bs, block_num), dtype=torch.int32)
    kv_cache = torch.randn(block_table.numel(), block_size, d).to(dtype).to(dev)
    ws = torch.empty(ops.cutlass_mla_get_workspace_size(block_num * block_size, bs, num_kv_splits=1), device=dev, dtype=torch.uint8)
    q_nope = torch.empty((num_heads, bs, dv), dtype=dtype, device=dev).transpose(0, 1)   # the reference's strided q_nope
    q_nope.copy_(q[:, :, :dv])
    q_pe = q[:, :, dv:].clone()
    out = ops.cutlass_mla_decode(q_nope, q_pe, kv_cache, seq_lens.to(dev), block_table.to(dev), ws, scale, 1)
    assert out.shape == (bs, num_heads, dv) and out.dtype == dtype
    ref = ref_mla(q, kv_cache, scale, block_table.to(dev), seq_lens)
    torch.testing.assert_close(out.float(), ref, atol=1e-2, rtol=1e-2)
    # the CPU oracle (exact softmax) over the token-granular view of the same page table
    if mean_seq_len <= 1024 and num_heads <= 32:
        r2t = page_table_to_token_slots(block_table, block_size, max_seq_len)
        rows = kv_cache.cpu().view(-1, 1, d)
        o = torch.zeros(bs, num_heads, dv, dtype=dtype)
        orc.decode_attention(q.cpu(), rows, rows.narrow(2, 0, dv), o, None, None, None, None, r2t, torch.arange(bs),
                             seq_lens.long(), scale, 0.0, is_mla=False)
        assert rel_l2(out, o) < 2e-3


def test_cutlass_mla_decode_asserts_like_the_reference():
    from iaas_sglang_b200 import ops

    q_nope = torch.zeros(2, 16, 512, dtype=torch.bfloat16, device="cuda")
    q_pe = torch.zeros(2, 16, 64, dtype=torch.bfloat16, device="cuda")
    cache = torch.zeros(4, 128, 576, dtype=torch.bfloat16, device="cuda")
    pt = torch.zeros(2, 2, dtype=torch.int32, device="cuda")
    ws = torch.empty(ops.cutlass_mla_get_workspace_size(256, 2), dtype=torch.uint8, device="cuda")
    with pytest.raises(AssertionError, match="int32"):
        ops.cutlass_mla_decode(q_nope, q_pe, cache, torch.ones(2, dtype=torch.int64, device="cuda"), pt, ws, 0.1)
    with pytest.raises(AssertionError):
        ops.cutlass_mla_decode(q_nope, q_pe[:, :8], cache, torch.ones(2, dtype=torch.int32, device="cuda"), pt, ws, 0.1)
    with pytest.raises(AssertionError, match="max_seq_len"):
        ops.cutlass_mla_get_workspace_size(0, 2)


def _mla_plugin(B, Hq, ctx, lens, meta_dtype=torch.int64):
    from iaas_sglang_b200 import compat
    from tests.test_gpu_backend import make_runner

    runner, backend = make_runner(B, Hq, 1, 576, ctx, layers=1, mla=True)
    c = DecodeCase(B, Hq, 1, 576, 512, lens, max_ctx=ctx, extra_reqs=3)
    pool, r2t = runner.token_to_kv_pool, runner.req_to_token_pool
    r2t.req_to_token.copy_(c.req_to_token)
    T = c.k_buf.shape[0]
    pool.kv_buffer[0][:T].copy_(c.k_buf)
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, B, req_pool_indices=c.req_pool_indices.to(meta_dtype).cuda(),
                             seq_lens=c.seq_lens.to(meta_dtype).cuda(), out_cache_loc=c.loc.cuda(),
                             seq_lens_sum=int(c.seq_lens.sum()), seq_lens_cpu=c.seq_lens, req_to_token_pool=r2t,
                             token_to_kv_pool=pool, attn_backend=backend)
    backend.init_forward_metadata(fb)
    layer = compat.RadixAttention(Hq, 576, c.sm_scale, 1, 0, v_head_dim=512)
    return c, backend, fb, layer, pool, T


@pytest.mark.parametrize("Hq", [16, 22, 128])
@pytest.mark.parametrize("meta_dtype", [torch.int64, torch.int32], ids=["i64", "i32"])
def test_plugin_forward_decode_with_q_rope_and_k_rope(Hq, meta_dtype):
    """cutlass_mla_backend.py:226-285 call shape: q = q_nope [B, H*512], q_rope [B, H, 64], k = latent [B,1,512],
    k_rope [B,1,64].  Same kernel, same data as the packed call -> bit-identical output; the pool row is the concat."""
    B, ctx = 4, 700
    c, backend, fb, layer, pool, T = _mla_plugin(B, Hq, ctx, [700, 33, 256, 1], meta_dtype)
    q = c.q.cuda()
    q_nope = q[:, :, :512].contiguous().view(B, Hq * 512)
    q_rope = q[:, :, 512:].contiguous()
    k_new = c.k_new.cuda()
    k_nope, k_rope = k_new[:, :, :512].contiguous(), k_new[:, :, 512:].contiguous()
    o = layer(q_nope, k_nope, k_nope, fb, q_rope=q_rope, k_rope=k_rope)
    ref, kb, _ = c.oracle(round_p=False)
    assert o.shape == (B, Hq * 512)
    assert rel_l2(o.view(B, Hq, 512), ref) < 2e-3
    assert torch.equal(pool.kv_buffer[0][:T].cpu(), kb), "set_mla_kv_buffer: pool row != concat(k_nope, k_rope)"
    # packed form on a fresh pool: bit-identical
    c2, backend2, fb2, layer2, pool2, _ = _mla_plugin(B, Hq, ctx, [700, 33, 256, 1], meta_dtype)
    o2 = layer2(q.view(B, Hq * 576), k_new, c.v_new.cuda(), fb2)
    assert torch.equal(o, o2)


def test_plugin_rejects_unknown_keywords_and_gqa_rope_arguments():
    c, backend, fb, layer, pool, T = _mla_plugin(2, 16, 128, [128, 5])
    q = c.q.cuda().view(2, -1)
    with pytest.raises(TypeError):
        backend.forward_decode(q, c.k_new.cuda(), c.v_new.cuda(), layer, fb, sinks=None)
    with pytest.raises(TypeError):
        backend.forward_extend(q, c.k_new.cuda(), c.v_new.cuda(), layer, fb, foo=1)


def test_cuda_graph_with_the_reference_runner_dtypes():
    """cuda_graph_runner.py:273-276 allocates seq_lens and req_pool_indices as int32 (schedule_batch.py:1490 does the
    same for idle batches); out_cache_loc stays int64.  Capture with dummy metadata, replay with real contents."""
    from iaas_sglang_b200 import compat
    from tests.test_gpu_backend import make_runner

    B, Hq, Hkv, D, ctx = 4, 32, 8, 128, 2048
    runner, backend = make_runner(B, Hq, Hkv, D, ctx, layers=1)
    backend.init_cuda_graph_state(max_bs=B, max_num_tokens=B)
    c = DecodeCase(B, Hq, Hkv, D, D, [2048, 5, 777, 1500], max_ctx=ctx, extra_reqs=3)
    pool, r2t = runner.token_to_kv_pool, runner.req_to_token_pool
    T = c.k_buf.shape[0]
    pool.k_buffer[0][:T].copy_(c.k_buf)
    pool.v_buffer[0][:T].copy_(c.v_buf)
    seq_lens = torch.full((B,), backend.get_cuda_graph_seq_len_fill_value(), dtype=torch.int32, device="cuda")
    rpi = torch.zeros(B, dtype=torch.int32, device="cuda")
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
        layer(q, k, v, fb)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            out.copy_(layer(q, k, v, fb))
    r2t.req_to_token.copy_(c.req_to_token)
    seq_lens.copy_(c.seq_lens)
    rpi.copy_(c.req_pool_indices)
    loc.copy_(c.loc)
    q.copy_(c.q.view(B, -1)); k.copy_(c.k_new.view(B, -1)); v.copy_(c.v_new.view(B, -1))
    backend.init_forward_metadata_replay_cuda_graph(B, rpi, seq_lens, int(c.seq_lens.sum()), None, compat.ForwardMode.DECODE, None,
                                                    seq_lens_cpu=c.seq_lens)
    graph.replay()
    torch.cuda.synchronize()
    ref, kb, vb = c.oracle()
    assert rel_l2(out.view(B, Hq, D), ref) < 2e-3
    assert torch.equal(pool.k_buffer[0][:T].cpu()[1:], kb[1:])
