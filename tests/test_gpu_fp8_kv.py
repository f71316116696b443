"""GPU tier: FP8 (e4m3) KV cache on the tuned decode path (memory_pool.py:116-121,431-440; kv_cache.py): the pool holds
codes, K = code * k_scale, V = code * v_scale.  Oracle = the reference CPU algorithm (oracle.cpu_oracle.decode_attention)
on the DEQUANTISED pool -- with power-of-two scales the dequantised values are exact in bf16, so the usual 2e-3 holds; with
arbitrary scales the check is a plain fp32 torch attention over the fp32 dequantised pool."""
import pytest
import torch

from tests.util import DecodeCase, rel_l2

pytestmark = pytest.mark.gpu
E4M3 = torch.float8_e4m3fn


def quantise(x, scale):
    return (x.float() / scale).to(E4M3)


@pytest.mark.parametrize("B,Hq,Hkv,lens", [(4, 32, 8, [700, 33, 1, 1500]), (3, 8, 1, [2000, 129, 640]), (5, 20, 4, [300, 301, 17, 999, 128]),
                                           (2, 16, 2, [4096, 77])])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
def test_fp8_kv_decode_against_oracle_on_the_dequantised_pool(B, Hq, Hkv, lens, dtype):
    from iaas_sglang_b200 import ops

    D, ks, vs = 128, 2.0 ** -4, 2.0 ** -5
    c = DecodeCase(B, Hq, Hkv, D, D, lens, dtype=dtype, extra_reqs=2, seed=B + Hq)
    kq, vq = quantise(c.k_buf, ks), quantise(c.v_buf, vs)
    knq, vnq = quantise(c.k_new, ks), quantise(c.v_new, vs)
    # the oracle sees the dequantised pool and the dequantised new rows (exact in bf16 / f16: power-of-two scales)
    c.k_buf, c.v_buf = (kq.float() * ks).to(dtype), (vq.float() * vs).to(dtype)
    k_new_raw, v_new_raw = c.k_new, c.v_new
    c.k_new, c.v_new = (knq.float() * ks).to(dtype), (vnq.float() * vs).to(dtype)
    ref, kb, vb = c.oracle()
    d = c.cuda()
    kpool, vpool = kq.cuda(), vq.cuda()
    out = torch.empty(B, Hq, D, dtype=dtype, device="cuda")
    ops.decode_attention(d["q"], kpool, vpool, out, k_new_raw.cuda(), v_new_raw.cuda(), d["loc"], None, d["req_to_token"],
                         d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0, k_scale=ks, v_scale=vs)
    assert rel_l2(out, ref) < 2e-3, rel_l2(out, ref)
    # the append wrote e4m3(x / scale): bit-exact against torch's cast
    loc = c.loc.long()
    assert torch.equal(kpool.cpu().view(torch.uint8)[loc], knq.view(torch.uint8))
    assert torch.equal(vpool.cpu().view(torch.uint8)[loc], vnq.view(torch.uint8))
    # uint8 storage view (what the reference pool hands out before .view(dtype)) and no append
    out2 = torch.empty_like(out)
    ops.decode_attention(d["q"], kpool.view(torch.uint8), vpool.view(torch.uint8), out2, None, None, None, None, d["req_to_token"],
                         d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0, k_scale=ks, v_scale=vs)
    assert torch.equal(out, out2)


def test_fp8_kv_arbitrary_scales_against_fp32_attention():
    from iaas_sglang_b200 import ops

    B, Hq, Hkv, D, ks, vs = 3, 16, 4, 128, 0.0371, 0.0219
    c = DecodeCase(B, Hq, Hkv, D, D, [900, 64, 333], seed=7)
    kq, vq = quantise(c.k_buf, ks), quantise(c.v_buf, vs)
    d = c.cuda()
    out = torch.empty(B, Hq, D, dtype=torch.bfloat16, device="cuda")
    ops.decode_attention(d["q"], kq.cuda(), vq.cuda(), out, None, None, None, None, d["req_to_token"], d["req_pool_indices"],
                         d["seq_lens"], c.sm_scale, 0.0, k_scale=ks, v_scale=vs)
    G = Hq // Hkv
    for b in range(B):
        L = int(c.seq_lens[b])
        slots = c.req_to_token[c.req_pool_indices[b], :L].long()
        k = (kq[slots].float() * ks).repeat_interleave(G, dim=1)
        v = (vq[slots].float() * vs).repeat_interleave(G, dim=1)
        s = torch.einsum("hd,lhd->hl", c.q[b].float(), k) * c.sm_scale
        ref = torch.einsum("hl,lhd->hd", torch.softmax(s, dim=-1), v)
        assert rel_l2(out[b], ref) < 2e-3


def test_fp8_kv_rejects_what_the_tuned_kernel_does_not_cover():
    from iaas_sglang_b200 import ops

    c = DecodeCase(2, 8, 2, 64, 64, [100, 50], seed=1)
    d = c.cuda()
    out = torch.empty(2, 8, 64, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(NotImplementedError, match="head_dim 128"):
        ops.decode_attention(d["q"], quantise(c.k_buf, 1.0).cuda(), quantise(c.v_buf, 1.0).cuda(), out, None, None, None, None,
                             d["req_to_token"], d["req_pool_indices"], d["seq_lens"], c.sm_scale, 0.0, k_scale=1.0, v_scale=1.0)


def test_plugin_with_an_fp8_pool():
    """compat.MHATokenToKVPool(dtype=float8_e4m3fn) stores uint8 like the reference pool; forward_decode passes the layer's
    k_scale / v_scale; forward_extend's set_kv_buffer quantises."""
    from iaas_sglang_b200 import compat
    from iaas_sglang_b200.backend import B200AttnBackend

    B, Hq, Hkv, D, ctx, ks, vs = 3, 16, 4, 128, 600, 2.0 ** -4, 2.0 ** -4
    c = DecodeCase(B, Hq, Hkv, D, D, [600, 100, 257], max_ctx=ctx, extra_reqs=3, seed=3)
    kq, vq = quantise(c.k_buf, ks), quantise(c.v_buf, vs)
    knq, vnq = quantise(c.k_new, ks), quantise(c.v_new, vs)
    k_new_raw, v_new_raw = c.k_new, c.v_new
    c.k_buf, c.v_buf = (kq.float() * ks).to(torch.bfloat16), (vq.float() * vs).to(torch.bfloat16)
    c.k_new, c.v_new = (knq.float() * ks).to(torch.bfloat16), (vnq.float() * vs).to(torch.bfloat16)
    ref, _, _ = c.oracle()
    dev = torch.device("cuda", 0)
    T = kq.shape[0]
    r2t = compat.ReqToTokenPool(B + 3, ctx, dev)
    kv = compat.MHATokenToKVPool(T, 1, E4M3, Hkv, D, 1, dev)
    assert kv.k_buffer[0].dtype == torch.uint8 and kv.get_key_buffer(0).dtype == E4M3
    kv.k_buffer[0][:T].copy_(kq.view(torch.uint8))
    kv.v_buffer[0][:T].copy_(vq.view(torch.uint8))
    r2t.req_to_token.copy_(c.req_to_token)
    runner = compat.ModelRunnerStub(compat.ModelConfigStub(Hq, Hkv, D, ctx), r2t, kv, device=dev)
    be = B200AttnBackend(runner)
    layer = compat.RadixAttention(Hq, D, c.sm_scale, Hkv, 0)
    layer.k_scale_float, layer.v_scale_float = ks, vs
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, B, req_pool_indices=c.req_pool_indices.cuda(), seq_lens=c.seq_lens.cuda(),
                             out_cache_loc=c.loc.cuda(), seq_lens_sum=int(c.seq_lens.sum()), seq_lens_cpu=c.seq_lens,
                             req_to_token_pool=r2t, token_to_kv_pool=kv, attn_backend=be)
    be.init_forward_metadata(fb)
    o = layer(c.q.cuda().view(B, -1), k_new_raw.cuda().view(B, -1), v_new_raw.cuda().view(B, -1), fb)
    assert rel_l2(o.view(B, Hq, D), ref) < 2e-3
    assert torch.equal(kv.k_buffer[0].cpu()[c.loc.long()], knq.view(torch.uint8))
    # the pool's own scatter (prefill path)
    kv.set_kv_buffer(layer, c.loc.cuda(), v_new_raw.cuda(), k_new_raw.cuda(), ks, vs)
    assert torch.equal(kv.k_buffer[0].cpu()[c.loc.long()], vnq.view(torch.uint8))
