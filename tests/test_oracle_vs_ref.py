"""CPU tier: the plain-C oracle against the reference's own CPU kernels (oracle/_ref) run LIVE, at the reference
tests' full shapes (test/srt/cpu/test_decode.py:146-158, test_mla.py:146-150).  Skipped where oracle/_ref cannot
run (no AVX512-BF16/AMX host, or the prebuilt library is absent); the committed golden vectors
(test_oracle_golden.py) cover the same ground everywhere."""
import pytest
import torch

from oracle import cpu_oracle as orc
from oracle import ref_ops
from tests.util import DecodeCase, max_abs, rel_l2

pytestmark = pytest.mark.skipif(not ref_ops.available(), reason=f"oracle/_ref unavailable: {ref_ops.why_unavailable()}")

DECODE_SHAPES = [  # test_decode.py:146-158
    (2, 16, 16, 64, 64), (2, 16, 1, 16, 16), (2, 32, 8, 33, 55), (2, 16, 1, 64, 64), (2, 64, 1, 13, 13),
    (2, 128, 1, 80, 80), (2, 128, 2, 512, 512), (1, 16, 1, 576, 512), (1, 16, 16, 576, 512), (1, 22, 1, 576, 512),
    (1, 40, 8, 128, 128),
]


def _ref_decode(c: DecodeCase, cap=0.0):
    kb = c.k_buf.clone()
    vb = kb.narrow(2, 0, c.Dv) if c.mla else c.v_buf.clone()
    out = torch.zeros(c.B, c.Hq, c.Dv, dtype=c.dtype)
    logits = torch.zeros(c.B, c.Hq, 8, c.Dv + 1, dtype=torch.float32)
    ref_ops.decode_attention_cpu(c.q, kb, vb, out, c.k_new, c.v_new, c.loc, logits, c.req_to_token, c.req_pool_indices,
                                 c.seq_lens, c.sm_scale, cap)
    return out, kb, vb


@pytest.mark.parametrize("shape", DECODE_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_decode_reference_shapes(shape):
    B, Hq, Hkv, D, Dv = shape
    c = DecodeCase(B, Hq, Hkv, D, Dv, 1024, mla=(Hkv == 1 and D == Dv + 64))
    ref, kb_r, vb_r = _ref_decode(c)
    out, kb, vb = c.oracle()
    assert rel_l2(out, ref) < 2e-3 and max_abs(out, ref) <= 2 ** -7
    assert torch.equal(kb, kb_r) and torch.equal(vb, vb_r)


@pytest.mark.parametrize("B,S", [(1, 8 * 111), (4, 8 * 128), (8, 8 * 133)])  # test_mla.py:146-150 (40 -> 8 requests)
def test_mla_reference_shapes(B, S):
    c = DecodeCase(B, 22, 1, 576, 512, S)
    ref, kb_r, _ = _ref_decode(c)
    out, kb, _ = c.oracle()
    assert rel_l2(out, ref) < 2e-3
    assert torch.equal(kb, kb_r)


def test_decode_variable_lengths_int64_table_and_cap():
    c = DecodeCase(5, 32, 8, 128, 128, [1024, 700, 333, 64, 57], idx_dtype=torch.int64, extra_reqs=3)
    for cap in (0.0, 30.0):
        ref, _, _ = _ref_decode(c, cap)
        out, _, _ = c.oracle(logit_cap=cap)
        assert rel_l2(out, ref) < 2e-3


def test_config1_shape_shuffled_page_table():
    """BASELINE.json configs[0]: Llama-3.2-1B-shaped, batch 16 x 1k ctx, page_size 1 (shuffled slots)."""
    c = DecodeCase(16, 32, 8, 64, 64, 1024)
    ref, kb_r, vb_r = _ref_decode(c)
    out, kb, vb = c.oracle()
    assert rel_l2(out, ref) < 2e-3
    assert torch.equal(kb, kb_r) and torch.equal(vb, vb_r)
