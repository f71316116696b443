"""GPU tier, collected last: b200_decode_attn_static (the entry point that allocates nothing: its split-KV scratch is a
file-scope device array inside the library) against b200_decode_attn with a caller-provided workspace, through the raw
C ABI, on two calls in a row (different batches, so stale scratch of the first call would show in the second), plus the
size limit.  Tolerance: the repo's attention bar (rel-L2 2e-3); the two entry points run the same kernels, so the outputs
are in fact expected to be equal."""
import ctypes as C

import pytest
import torch

from tests.util import rel_l2

pytestmark = pytest.mark.gpu


def _params(_lib, q, k_buf, v_buf, out, r2t, seq_lens, scale, num_splits=0):
    B, Hq, D = q.shape
    p = _lib.DecodeAttnParams()
    p.q, p.q_stride_b, p.q_stride_h = q.data_ptr(), q.stride(0), q.stride(1)
    p.k_buf, p.kbuf_stride_t, p.kbuf_stride_h = k_buf.data_ptr(), k_buf.stride(0), k_buf.stride(1)
    p.v_buf, p.vbuf_stride_t, p.vbuf_stride_h = v_buf.data_ptr(), v_buf.stride(0), v_buf.stride(1)
    p.out = out.data_ptr()
    p.req_to_token, p.req_to_token_stride = r2t.data_ptr(), r2t.stride(0)
    p.idx_is_i64 = int(r2t.dtype == torch.int64)
    p.req_pool_indices = None            # request b owns row b of the table
    p.seq_lens = seq_lens.data_ptr()
    p.meta_i32_mask = int(seq_lens.dtype == torch.int32)
    p.page_size = 1
    p.B, p.Hq, p.Hkv, p.D, p.Dv = B, Hq, k_buf.shape[1], D, v_buf.shape[2]
    p.dtype = _lib.DTYPE_CODE[q.dtype]
    p.write_kv = 0
    p.max_seq_len = int(r2t.shape[1])
    p.num_splits = num_splits
    p.sm_scale, p.logit_cap = float(scale), 0.0
    return p


def _case(B, Hq, Hkv, D, lens, seed):
    g = torch.Generator().manual_seed(seed)
    T = int(sum(lens))
    k_buf = torch.randn(T + 1, Hkv, D, generator=g).to(torch.bfloat16).cuda()
    v_buf = torch.randn(T + 1, Hkv, D, generator=g).to(torch.bfloat16).cuda()
    q = torch.randn(B, Hq, D, generator=g).to(torch.bfloat16).cuda()
    slots = (torch.randperm(T, generator=g) + 1).to(torch.int32)
    r2t = torch.zeros(B, max(lens), dtype=torch.int32)
    off = 0
    for b, n in enumerate(lens):
        r2t[b, :n] = slots[off: off + n]
        off += n
    return q, k_buf, v_buf, r2t.cuda(), torch.tensor(lens, dtype=torch.int64).cuda()


# (B, Hq, Hkv, D, seq lens, split hint): the tuned GQA kernel (split partials in the scratch), then the generic kernel
CALLS = [
    (5, 32, 8, 128, [700, 33, 1500, 1, 260], 0),
    (3, 16, 2, 128, [2100, 64, 515], 0),
    (4, 8, 2, 96, [900, 300, 17, 1024], 4),
]


def test_static_scratch_entry_matches_workspace_entry_twice():
    from iaas_sglang_b200 import _lib

    lib = _lib.load()
    stream = _lib.current_stream()
    for rep in range(2):                 # the whole sequence twice: the array keeps the previous call's contents
        for i, (B, Hq, Hkv, D, lens, hint) in enumerate(CALLS):
            q, k_buf, v_buf, r2t, sl = _case(B, Hq, Hkv, D, lens, seed=10 * rep + i)
            want = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device="cuda")
            got = torch.zeros_like(want)
            p = _params(_lib, q, k_buf, v_buf, want, r2t, sl, D ** -0.5, hint)
            S = lib.b200_decode_attn_num_splits(B, Hq, Hkv, D, D, p.max_seq_len, hint)
            p.num_splits = S
            need = lib.b200_decode_attn_workspace_size(B, Hq, D, S)
            ws = torch.empty(need, dtype=torch.uint8, device="cuda")
            p.workspace, p.workspace_bytes = ws.data_ptr(), need
            _lib.check(lib.b200_decode_attn(p, stream))
            ps = _params(_lib, q, k_buf, v_buf, got, r2t, sl, D ** -0.5, hint)
            ps.num_splits = S            # workspace / workspace_bytes stay NULL / 0: the entry point must not need them
            _lib.check(lib.b200_decode_attn_static(ps, stream))
            torch.cuda.synchronize()
            assert torch.isfinite(got.float()).all()
            assert rel_l2(got, want) < 2e-3, (rep, i, rel_l2(got, want))
            assert torch.equal(got, want), (rep, i)


def test_static_scratch_entry_rejects_sizes_beyond_the_array():
    from iaas_sglang_b200 import _lib

    lib = _lib.load()
    p = _lib.DecodeAttnParams()          # sizes only: the check comes before any pointer is looked at
    p.B, p.Hq, p.Hkv, p.D, p.Dv = 4096, 32, 8, 96, 96
    p.num_splits, p.max_seq_len = 16, 8192
    p.dtype = _lib.DTYPE_CODE[torch.bfloat16]
    rc = lib.b200_decode_attn_static(p, None)
    assert rc == _lib.B200_ERR_ARG
    msg = _lib.last_error()
    assert "static array holds" in msg and "decode_attn_static" in msg
