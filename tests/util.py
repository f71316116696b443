"""Shared builders for the parity tests: seeded inputs in the reference's own test distributions
(test/srt/cpu/test_decode.py:62-143, sgl-kernel/tests/test_fp8_blockwise_gemm.py:62-81, ...)."""
from __future__ import annotations

import torch

from oracle import cpu_oracle as orc


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def max_abs(a: torch.Tensor, b: torch.Tensor) -> float:
    return (a.float().cpu() - b.float().cpu()).abs().max().item()


class DecodeCase:
    """One paged decode-attention problem on CPU tensors (oracle side); .cuda() gives the device copy."""

    def __init__(self, B, Hq, Hkv, D, Dv, seq_lens, *, dtype=torch.bfloat16, seed=1234, shuffle=True, mla=None,
                 idx_dtype=torch.int32, max_ctx=None, shared_prefix=0, extra_reqs=0, noncontig=False):
        g = torch.Generator().manual_seed(seed)
        if isinstance(seq_lens, int):
            seq_lens = [seq_lens] * B
        self.B, self.Hq, self.Hkv, self.D, self.Dv = B, Hq, Hkv, D, Dv
        self.dtype = dtype
        self.mla = (Hkv == 1 and D == Dv + 64) if mla is None else mla
        sl = torch.tensor(seq_lens, dtype=torch.int64)
        max_ctx = max_ctx or int(sl.max())
        n_unique = int(sl.sum()) - shared_prefix * (B - 1) if shared_prefix else int(sl.sum())
        T = n_unique + 1  # + reserved slot 0
        self.q = torch.randn(B, Hq, D, generator=g).to(dtype)
        if noncontig:  # head-major storage viewed as [T,H,D] (test_decode.py:106-110)
            self.k_buf = torch.randn(Hkv, T, D, generator=g).to(dtype).transpose(0, 1)
        else:
            self.k_buf = torch.randn(T, Hkv, D, generator=g).to(dtype)
        if self.mla:
            self.v_buf = self.k_buf.narrow(2, 0, Dv)
        elif noncontig:
            self.v_buf = torch.randn(Hkv, T, Dv, generator=g).to(dtype).transpose(0, 1)
        else:
            self.v_buf = torch.randn(T, Hkv, Dv, generator=g).to(dtype)
        self.k_new = torch.randn(B, Hkv, D, generator=g).to(dtype)
        self.v_new = self.k_new.narrow(2, 0, Dv) if self.mla else torch.randn(B, Hkv, Dv, generator=g).to(dtype)
        slots = (torch.randperm(n_unique, generator=g) + 1) if shuffle else (torch.arange(n_unique) + 1)
        n_req = B + extra_reqs
        r2t = torch.zeros(n_req, max_ctx, dtype=torch.int64)
        rpi = torch.randperm(n_req, generator=g)[:B] if extra_reqs else torch.arange(B)
        off = 0
        shared = None
        if shared_prefix:
            shared = slots[:shared_prefix]
            off = shared_prefix
        for b in range(B):
            L = int(sl[b])
            if shared_prefix:
                r2t[rpi[b], :shared_prefix] = shared
                r2t[rpi[b], shared_prefix:L] = slots[off:off + L - shared_prefix]
                off += L - shared_prefix
            else:
                r2t[rpi[b], :L] = slots[off:off + L]
                off += L
        self.req_to_token = r2t.to(idx_dtype)
        self.req_pool_indices = rpi.to(torch.int64)
        self.seq_lens = sl
        self.loc = torch.stack([r2t[rpi[b], int(sl[b]) - 1] for b in range(B)]).to(torch.int64)
        self.sm_scale = 1.0 / (D ** 0.5)

    def oracle(self, logit_cap=0.0, write_kv=True, num_kv_splits=8, round_p=None):
        """round_p: the reference's MLA path rounds the softmax weights to bf16 before P.V (decode.cpp:1157-1159);
        None = as the reference does (MLA only), False = keep them in fp32 (the exact softmax)."""
        kb = self.k_buf.clone()
        vb = kb.narrow(2, 0, self.Dv) if self.mla else self.v_buf.clone()
        out = torch.zeros(self.B, self.Hq, self.Dv, dtype=self.dtype)
        orc.decode_attention(self.q, kb, vb, out, self.k_new if write_kv else None, self.v_new if write_kv else None,
                             self.loc, None, self.req_to_token, self.req_pool_indices, self.seq_lens, self.sm_scale,
                             logit_cap, num_kv_splits=num_kv_splits, is_mla=self.mla if round_p is None else round_p)
        return out, kb, vb

    def cuda(self, device="cuda"):
        d = {}
        kb = self.k_buf.to(device)
        if self.k_buf.stride(1) != self.D:  # keep the non-contiguous layout on the device too
            kb = self.k_buf.transpose(0, 1).contiguous().to(device).transpose(0, 1)
        d["k_buf"] = kb
        if self.mla:
            d["v_buf"] = kb.narrow(2, 0, self.Dv)
        elif self.v_buf.stride(1) != self.Dv:
            d["v_buf"] = self.v_buf.transpose(0, 1).contiguous().to(device).transpose(0, 1)
        else:
            d["v_buf"] = self.v_buf.to(device)
        d["q"] = self.q.to(device)
        d["k_new"] = self.k_new.to(device)
        d["v_new"] = d["k_new"].narrow(2, 0, self.Dv) if self.mla else self.v_new.to(device)
        for n in ("req_to_token", "req_pool_indices", "seq_lens", "loc"):
            d[n] = getattr(self, n).to(device)
        return d


def run_cuda_decode(case: DecodeCase, logit_cap=0.0, write_kv=True, fused=False, num_splits=0, device="cuda"):
    from iaas_sglang_b200 import ops

    d = case.cuda(device)
    out = torch.zeros(case.B, case.Hq, case.Dv, dtype=case.dtype, device=device)
    ops.decode_attention(d["q"], d["k_buf"], d["v_buf"], out, d["k_new"] if write_kv else None,
                         d["v_new"] if write_kv else None, d["loc"], None, d["req_to_token"], d["req_pool_indices"],
                         d["seq_lens"], case.sm_scale, logit_cap, max_seq_len=int(case.seq_lens.max()),
                         num_splits=num_splits, fused_kv_write=fused)
    torch.cuda.synchronize()
    return out.cpu(), d["k_buf"].cpu(), d["v_buf"].cpu()


def make_fp8_gemm_case(M, N, K, seed=0, out_dtype=torch.bfloat16, dist="uniform448"):
    """Inputs of sgl-kernel/tests/test_fp8_blockwise_gemm.py:62-81 (uniform in +-448, scales randn*1e-3)."""
    g = torch.Generator().manual_seed(seed)
    if dist == "uniform448":
        a32 = (torch.rand(M, K, generator=g) - 0.5) * 2 * 448.0
        b32 = (torch.rand(N, K, generator=g) - 0.5) * 2 * 448.0
        sa = torch.randn(M, (K + 127) // 128, generator=g) * 0.001
        sb = torch.randn((K + 127) // 128, (N + 127) // 128, generator=g) * 0.001
    else:  # realistic: quantised activations/weights
        a32 = torch.randn(M, K, generator=g) * 40
        b32 = torch.randn(N, K, generator=g) * 40
        sa = torch.rand(M, (K + 127) // 128, generator=g) * 0.01 + 0.001
        sb = torch.rand((K + 127) // 128, (N + 127) // 128, generator=g) * 0.01 + 0.001
    a = a32.clamp(-448, 448).to(torch.float8_e4m3fn)
    b = b32.clamp(-448, 448).to(torch.float8_e4m3fn).t()  # [K,N] column-major
    sa = sa.t().contiguous().t()
    sb = sb.t().contiguous().t()
    return a, b, sa, sb


# ---------------------------------------------------------------------------------------------
# golden fixtures (tests/golden/*.npz, written by tests/golden/make_golden.py from the reference itself)
# ---------------------------------------------------------------------------------------------
import os as _os

import numpy as _np

GOLDEN_DIR = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden")


def golden(name: str):
    return _np.load(_os.path.join(GOLDEN_DIR, name + ".npz"))


def bf16(a) -> torch.Tensor:
    return torch.from_numpy(_np.ascontiguousarray(a).view(_np.int16)).view(torch.bfloat16)


def f16(a) -> torch.Tensor:
    return torch.from_numpy(_np.ascontiguousarray(a).view(_np.int16)).view(torch.float16)


def fp8(a) -> torch.Tensor:
    return torch.from_numpy(_np.ascontiguousarray(a)).view(torch.float8_e4m3fn)


def golden_decode_cases():
    """Yield dicts of CPU tensors for every case of tests/golden/decode_attention_cpu.npz."""
    g = golden("decode_attention_cpu")
    for i in range(int(g["n_cases"])):
        p = f"c{i}_"
        B, Hq, Hkv, D, Dv, S, mla = [int(v) for v in g[p + "shape"]]
        k_buf = bf16(g[p + "k_buf"])
        key = bf16(g[p + "key"])
        c = dict(idx=i, B=B, Hq=Hq, Hkv=Hkv, D=D, Dv=Dv, S=S, mla=bool(mla), cap=float(g[p + "cap"]), q=bf16(g[p + "q"]),
                 k_buf=k_buf, key=key, r2t=torch.from_numpy(g[p + "r2t"]), seq_lens=torch.from_numpy(g[p + "seq_lens"]),
                 loc=torch.from_numpy(g[p + "loc"]), o=bf16(g[p + "o"]), rpi=torch.arange(B, dtype=torch.int64))
        if mla:
            c["v_buf"], c["value"] = k_buf.narrow(2, 0, Dv), key.narrow(2, 0, Dv)
        else:
            c["v_buf"], c["value"] = bf16(g[p + "v_buf"]), bf16(g[p + "value"])
        yield c


def golden_extend_cases():
    g = golden("extend_attention_cpu")
    for i in range(int(g["n_cases"])):
        p = f"c{i}_"
        B, Hq, Hkv, D, Dv = [int(v) for v in g[p + "shape"]]
        k_buf, v_buf = bf16(g[p + "k_buf"]), bf16(g[p + "v_buf"])
        r2t = torch.from_numpy(g[p + "r2t"])
        sl, el, es = (torch.from_numpy(g[p + n]) for n in ("seq_lens", "ext_lens", "ext_start"))
        Te = int(el.sum())
        k_ext = torch.empty(Te, Hkv, D, dtype=torch.bfloat16)
        v_ext = torch.empty(Te, Hkv, Dv, dtype=torch.bfloat16)
        for b in range(B):
            slots = r2t[b, int(sl[b] - el[b]): int(sl[b])].long()
            k_ext[int(es[b]): int(es[b] + el[b])] = k_buf[slots]
            v_ext[int(es[b]): int(es[b] + el[b])] = v_buf[slots]
        yield dict(idx=i, B=B, Hq=Hq, Hkv=Hkv, D=D, Dv=Dv, q=bf16(g[p + "q"]), k_buf=k_buf, v_buf=v_buf, k_ext=k_ext,
                   v_ext=v_ext, r2t=r2t, seq_lens=sl, ext_lens=el, ext_start=es, rpi=torch.arange(B, dtype=torch.int64),
                   o=bf16(g[p + "o"]))
