"""GPU tier: the tuned (tiled, warp tensor core) extend / prefill attention kernel against the CPU oracle
(oracle.cpu_oracle.extend_attention = the restatement of sgl-kernel/csrc/cpu/extend.cpp) on seeded ragged batches, at the
reference's own tolerance (test/srt/cpu/test_extend.py:183, 1e-2), plus the two extras the cascade path uses
(prefix_only, lse_out) against a plain fp32 torch attention."""
import pytest
import torch

from oracle import cpu_oracle as orc
from tests.util import rel_l2

pytestmark = pytest.mark.gpu


def make_case(B, Hq, Hkv, D, prefix_lens, ext_lens, seed=0, dtype=torch.bfloat16, idx_dtype=torch.int32):
    g = torch.Generator().manual_seed(seed)
    pre = torch.tensor(prefix_lens, dtype=torch.int64)
    ext = torch.tensor(ext_lens, dtype=torch.int64)
    sl = pre + ext
    T = int(sl.sum())
    max_ctx = int(sl.max())
    k_buf = torch.randn(T + 1, Hkv, D, generator=g).to(dtype)
    v_buf = torch.randn(T + 1, Hkv, D, generator=g).to(dtype)
    slots = torch.randperm(T, generator=g) + 1
    r2t = torch.zeros(B, max_ctx, dtype=idx_dtype)
    off = 0
    for b in range(B):
        r2t[b, : int(sl[b])] = slots[off: off + int(sl[b])].to(idx_dtype)
        off += int(sl[b])
    es = torch.cumsum(ext, 0) - ext
    Te = int(ext.sum())
    q = torch.randn(Te, Hq, D, generator=g).to(dtype)
    k_ext = torch.empty(Te, Hkv, D, dtype=dtype)
    v_ext = torch.empty(Te, Hkv, D, dtype=dtype)
    for b in range(B):
        s_ = r2t[b, int(pre[b]): int(sl[b])].long()
        k_ext[int(es[b]): int(es[b] + ext[b])] = k_buf[s_]
        v_ext[int(es[b]): int(es[b] + ext[b])] = v_buf[s_]
    return dict(q=q, k_ext=k_ext, v_ext=v_ext, k_buf=k_buf, v_buf=v_buf, r2t=r2t, rpi=torch.arange(B, dtype=torch.int64),
                seq_lens=sl, ext_lens=ext, ext_start=es, pre=pre, scale=D ** -0.5)


CASES = [
    # B, Hq, Hkv, D, prefix lens, extend lens
    (1, 8, 8, 128, [0], [200]),                       # MHA, no prefix, one tile boundary inside
    (3, 32, 8, 128, [0, 100, 513], [300, 1, 77]),     # Llama GQA 4, ragged, a single-token extend
    (2, 16, 2, 64, [64, 0], [129, 640]),              # D 64, GQA 8
    (4, 8, 1, 128, [1000, 7, 0, 333], [5, 63, 64, 65]),   # TP8-local 70B shape: G 8, rows = tokens x 8
    (2, 40, 8, 128, [31, 2048], [1024, 96]),          # Qwen G 5 (rows of a tile straddle tokens unevenly)
]


@pytest.mark.parametrize("B,Hq,Hkv,D,pre,ext", CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
def test_tuned_extend_against_oracle(B, Hq, Hkv, D, pre, ext, dtype):
    from iaas_sglang_b200 import ops

    c = make_case(B, Hq, Hkv, D, pre, ext, seed=B + Hq + D, dtype=dtype)
    ref = torch.zeros(c["q"].shape[0], Hq, D, dtype=dtype)
    orc.extend_attention(c["q"], c["k_ext"], c["v_ext"], ref, c["k_buf"], c["v_buf"], c["r2t"], c["rpi"], c["seq_lens"],
                         c["ext_lens"], c["ext_start"], int(c["ext_lens"].max()), c["scale"], 0.0)
    d = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in c.items()}
    out = torch.zeros_like(d["q"])
    ops.extend_attention(d["q"], d["k_ext"], d["v_ext"], out, d["k_buf"], d["v_buf"], d["r2t"], d["rpi"], d["seq_lens"],
                         d["ext_lens"], d["ext_start"], int(c["ext_lens"].max()), c["scale"], 0.0)
    torch.testing.assert_close(out.float().cpu(), ref.float(), atol=1e-2, rtol=1e-2)   # test_extend.py:183
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)


def test_tuned_extend_logit_cap_and_int64_table():
    from iaas_sglang_b200 import ops

    c = make_case(2, 8, 2, 128, [50, 0], [70, 130], seed=9, idx_dtype=torch.int64)
    ref = torch.zeros(c["q"].shape[0], 8, 128, dtype=torch.bfloat16)
    orc.extend_attention(c["q"], c["k_ext"], c["v_ext"], ref, c["k_buf"], c["v_buf"], c["r2t"], c["rpi"], c["seq_lens"],
                         c["ext_lens"], c["ext_start"], 130, c["scale"], 30.0)
    d = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in c.items()}
    out = torch.zeros_like(d["q"])
    ops.extend_attention(d["q"], d["k_ext"], d["v_ext"], out, d["k_buf"], d["v_buf"], d["r2t"], d["rpi"], d["seq_lens"],
                         d["ext_lens"], d["ext_start"], 130, c["scale"], 30.0)
    assert rel_l2(out, ref) < 5e-3


def test_prefix_only_with_lse_against_fp32_attention():
    """prefix_only: row (token i, head h) = softmax(scale q k_prefix^T) v_prefix over exactly the request's prefix, and
    lse_out = natural-log log-sum-exp of the scaled logits (what merge_state consumes)."""
    from iaas_sglang_b200 import ops

    B, Hq, Hkv, D = 3, 16, 2, 128
    c = make_case(B, Hq, Hkv, D, [300, 0, 65], [9, 4, 130], seed=4)
    d = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in c.items()}
    Te = c["q"].shape[0]
    out = torch.zeros_like(d["q"])
    lse = torch.zeros(Te, Hq, dtype=torch.float32, device="cuda")
    ops.extend_attention(d["q"], d["k_ext"], d["v_ext"], out, d["k_buf"], d["v_buf"], d["r2t"], d["rpi"], d["seq_lens"],
                         d["ext_lens"], d["ext_start"], 130, c["scale"], 0.0, prefix_only=True, lse_out=lse)
    G = Hq // Hkv
    for b in range(B):
        P, e0, n = int(c["pre"][b]), int(c["ext_start"][b]), int(c["ext_lens"][b])
        o_b, l_b = out[e0:e0 + n].float().cpu(), lse[e0:e0 + n].cpu()
        if P == 0:
            assert torch.all(o_b == 0) and torch.all(torch.isinf(l_b) & (l_b < 0))
            continue
        slots = c["r2t"][b, :P].long()
        k = c["k_buf"][slots].float().repeat_interleave(G, dim=1)      # [P, Hq, D]
        v = c["v_buf"][slots].float().repeat_interleave(G, dim=1)
        s = torch.einsum("thd,phd->thp", c["q"][e0:e0 + n].float(), k) * c["scale"]
        ref_lse = torch.logsumexp(s, dim=-1)
        ref_o = torch.einsum("thp,phd->thd", torch.softmax(s, dim=-1), v)
        assert rel_l2(o_b, ref_o) < 5e-3
        torch.testing.assert_close(l_b, ref_lse, atol=2e-3, rtol=1e-4)
