#!/usr/bin/env python
"""Per-kernel roofline micro-benchmarks (developer tool; bench.py is the contract).

    python tools/kbench.py [attn] [gemm] [gemm_big] [int8] [awq] [mla] [elem]

Every case is timed with CUDA events over back-to-back launches after warm-up; operands rotate over a
set of buffers larger than the 126 MB L2, so each launch streams from HBM.  Prints one line per case:
algorithmic bytes / flops, time, GB/s and TFLOP/s, fraction of the measured peaks.
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from iaas_sglang_b200 import ops  # noqa: E402

DEV = torch.device("cuda", 0)
L2_BYTES = 160 << 20


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d["bf16_tflops"])
    return 6650.0, 1590.0


HBM, BF16_TF = peaks()


def timeit(fn, n_variants: int, reps: int = 0, warm: int = 3):
    """fn(i) launches variant i; returns ms per launch.  The launches are captured into one CUDA graph so that the
    host-side cost of the Python wrappers does not bound short kernels."""
    for i in range(max(warm, n_variants)):
        fn(i % n_variants)
    torch.cuda.synchronize()
    if reps == 0:
        reps = max(20, 2 * n_variants)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        fn(0)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for i in range(reps):
                fn(i % n_variants)
    torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def report(name, ms, nbytes, flops, tensor_peak_tf=None):
    gbs = nbytes / ms / 1e6
    tf = flops / ms / 1e9
    s = f"{name:<58s} {ms * 1e3:9.1f} us  {gbs:8.1f} GB/s ({gbs / HBM:5.1%} of measured HBM)  {tf:8.1f} TFLOP/s"
    if tensor_peak_tf:
        s += f" ({tf / tensor_peak_tf:5.1%} of {tensor_peak_tf:.0f})"
    print(s, flush=True)


# ------------------------------------------------------------------------------------------------
def attn_case(name, B, Hq, Hkv, D, ctx, shared=0, layers=2, fused=True, var=False, idx_dtype=torch.int32):
    g = torch.Generator(device=DEV).manual_seed(1234)
    unique = B * (ctx - shared) + shared
    perm = torch.randperm(unique, device=DEV, generator=g) + 1
    r2t = torch.zeros(B, ctx, dtype=idx_dtype, device=DEV)
    if shared:
        r2t[:, :shared] = perm[:shared].to(idx_dtype)[None]
    r2t[:, shared:] = perm[shared:].view(B, ctx - shared).to(idx_dtype)
    if var:
        seq = torch.randint(ctx // 2, ctx + 1, (B,), device=DEV, generator=g, dtype=torch.int64)
    else:
        seq = torch.full((B,), ctx, dtype=torch.int64, device=DEV)
    rpi = torch.arange(B, dtype=torch.int64, device=DEV)
    loc = r2t.gather(1, (seq - 1)[:, None]).view(-1).to(torch.int64)
    kbufs = [torch.randn(unique + 1, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
    vbufs = [torch.randn(unique + 1, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
    q = torch.randn(B, Hq, D, device=DEV, dtype=torch.bfloat16, generator=g)
    k = torch.randn(B, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g)
    v = torch.randn(B, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g)
    o = torch.empty(B, Hq, D, device=DEV, dtype=torch.bfloat16)
    S = ops.decode_attention_num_splits(B, Hq, Hkv, D, D, ctx)
    ws = torch.empty(ops.decode_attention_workspace_bytes(B, Hq, D, S), dtype=torch.uint8, device=DEV)

    def run(i):
        ops.decode_attention(q, kbufs[i], vbufs[i], o, k, v, loc, ws, r2t, rpi, seq, D ** -0.5, 0.0, max_seq_len=ctx,
                             fused_kv_write=fused)

    ms = timeit(run, layers, reps=20)
    tot = int(seq.sum())
    nbytes = tot * Hkv * 2 * D * 2 + tot * r2t.element_size() + B * Hq * 2 * D * 2 + B * Hkv * 2 * D * 2 * 2
    report(f"attn {name} B{B} H{Hq}/{Hkv} D{D} ctx{ctx} S{S}" + (" var" if var else ""), ms, nbytes, 2 * tot * Hq * 2 * D)
    del kbufs, vbufs
    torch.cuda.empty_cache()


def bench_attn():
    attn_case("cfg2", 64, 32, 8, 128, 4096)
    attn_case("cfg2", 64, 32, 8, 128, 4096, var=True)
    attn_case("cfg3tp2", 128, 20, 4, 128, 4096)
    attn_case("cfg3tp4", 128, 10, 2, 128, 4096)
    attn_case("cfg4tp8", 256, 8, 1, 128, 8192, shared=2048)
    attn_case("cfg1", 16, 32, 8, 64, 1024, layers=16)
    attn_case("small", 8, 32, 8, 128, 2048, layers=8)


# ------------------------------------------------------------------------------------------------
def fp8_case(M, N, K, tensor_peak=None):
    g = torch.Generator(device=DEV).manual_seed(0)
    nv = max(1, min(16, L2_BYTES // (N * K) + 1))
    ws_ = [(torch.randn(N, K, device=DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn) for _ in range(nv)]
    sb = (torch.rand((N + 127) // 128, (K + 127) // 128, device=DEV, generator=g) * 1e-3 + 1e-4)
    a = (torch.randn(M, K, device=DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sa = (torch.rand((K + 127) // 128, M, device=DEV, generator=g) * 1e-3 + 1e-4).t()
    sbt = sb.t().contiguous().t() if False else sb.t()   # [K/128, N/128] K-major view of the [N/128, K/128] tensor

    def run(i):
        ops.fp8_blockwise_scaled_mm(a, ws_[i].t(), sa, sbt, torch.bfloat16)

    ms = timeit(run, nv)
    nbytes = N * K + (N // 128) * (K // 128) * 4 + M * K + M * (K // 128) * 4 + M * N * 2
    report(f"fp8_blockwise M{M} N{N} K{K}", ms, nbytes, 2 * M * N * K, tensor_peak)


def int8_case(M, N, K, tensor_peak=None):
    g = torch.Generator(device=DEV).manual_seed(0)
    nv = max(1, min(16, L2_BYTES // (N * K) + 1))
    ws_ = [torch.randint(-127, 128, (N, K), device=DEV, generator=g, dtype=torch.int8) for _ in range(nv)]
    a = torch.randint(-127, 128, (M, K), device=DEV, generator=g, dtype=torch.int8)
    sa = torch.rand(M, device=DEV, generator=g) * 1e-3
    sb = torch.rand(N, device=DEV, generator=g) * 1e-3

    def run(i):
        ops.int8_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16, None)

    ms = timeit(run, nv)
    nbytes = N * K + N * 4 + M * K + M * 4 + M * N * 2
    report(f"int8_scaled_mm M{M} N{N} K{K}", ms, nbytes, 2 * M * N * K, tensor_peak)


def awq_case(M, N, K):
    g = torch.Generator(device=DEV).manual_seed(0)
    nv = max(1, min(16, L2_BYTES // (N * K // 2) + 1))
    qw = [torch.randint(0, 2 ** 31 - 1, (K, N // 8), device=DEV, dtype=torch.int32, generator=g) for _ in range(nv)]
    qz = torch.randint(0, 2 ** 31 - 1, (K // 128, N // 8), device=DEV, dtype=torch.int32, generator=g)
    sc = (torch.rand(K // 128, N, device=DEV, generator=g) * 4e-3 + 1e-3).to(torch.bfloat16)
    x = torch.randn(M, K, device=DEV, dtype=torch.bfloat16, generator=g)

    def run(i):
        ops.awq_w4a16_gemm(x, qw[i], sc, qz)

    ms = timeit(run, nv)
    nbytes = N * K // 2 + (K // 128) * N * 2 + (K // 128) * (N // 8) * 4 + M * K * 2 + M * N * 2
    report(f"awq_w4a16 M{M} N{N} K{K}", ms, nbytes, 2 * M * N * K)


CFG2_NK = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]
CFG4_NK = [(1280, 8192), (8192, 1024), (7168, 8192), (8192, 3584)]
CFG3_NK = [(3584, 5120), (5120, 2560), (27648, 5120), (5120, 13824)]


def bench_gemm():
    for n, k in CFG2_NK:
        fp8_case(64, n, k)
    for n, k in CFG4_NK:
        fp8_case(256, n, k)


def bench_gemm_big():
    pk = 2 * BF16_TF   # dense fp8 is twice the bf16 rate; both quoted: measured-bf16 x 2 and nominal 4500
    for m in (1024, 4096, 8192):
        fp8_case(m, 8192, 8192, pk)
    fp8_case(4096, 4096, 4096, pk)
    fp8_case(4096, 28672, 4096, pk)


def bench_int8():
    for n, k in CFG2_NK:
        int8_case(64, n, k)
    pk = 2 * BF16_TF
    for m in (1024, 4096, 8192):
        int8_case(m, 8192, 8192, pk)


def bench_awq():
    for n, k in CFG3_NK:
        awq_case(128, n, k)


def bench_prof():
    """few plain launches of the two attention shapes for ncu (-k regex:decode_attn_mma)"""
    global timeit

    def timeit(fn, n_variants, reps=0, warm=3):  # noqa: F811  (no graph, 4 launches)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(0); fn(1 % n_variants)
        e0.record()
        fn(0); fn(1 % n_variants)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 2

    attn_case("cfg2", 64, 32, 8, 128, 4096)
    attn_case("cfg3tp4", 128, 10, 2, 128, 4096)


def bench_proftp8():
    """a few plain launches of the TP8-local attention shapes (Hkv 1: 256-byte rows) for ncu (-k regex:decode_attn_mma)"""
    _plain_timeit()
    attn_case("cfg2tp8", 512, 4, 1, 128, 4096)


def bench_proftp24():
    """a few plain launches of the TP2- / TP4-local attention shapes of cfg 2 for ncu (-k regex:decode_attn_mma)"""
    _plain_timeit()
    attn_case("cfg2tp2", 128, 16, 4, 128, 4096)
    attn_case("cfg2tp4", 256, 8, 2, 128, 4096)


def bench_profgemm():
    """a few plain launches of decode-shape GEMMs for ncu (-k regex:gemm_skinny)"""
    global timeit

    def timeit(fn, n_variants, reps=0, warm=3):  # noqa: F811
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(0); fn(1 % n_variants)
        e0.record()
        fn(2 % n_variants); fn(3 % n_variants)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 2

    fp8_case(64, 4096, 14336)   # down_proj: 32 tiles x 8 K-slices
    fp8_case(64, 6144, 4096)    # qkv: 48 tiles x 5
    int8_case(64, 28672, 4096)


def bench_profbig():
    """a few plain launches of the large-M fp8 GEMM for ncu (-k regex:gemm_big)"""
    global timeit

    def timeit(fn, n_variants, reps=0, warm=3):  # noqa: F811
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(0); fn(1 % n_variants)
        e0.record()
        fn(0); fn(1 % n_variants)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 2

    fp8_case(4096, 8192, 8192, 2 * BF16_TF)


def mla_case(name, B, Hq, ctx, layers=2, var=False, splits=0):
    """MLA absorbed decode: latent pool [T, 1, 576], V = first 512 columns (cutlass_mla_decode's shapes, cfg 5)."""
    g = torch.Generator(device=DEV).manual_seed(1234)
    D, Dv = 576, 512
    unique = B * ctx
    perm = torch.randperm(unique, device=DEV, generator=g) + 1
    r2t = perm.view(B, ctx).to(torch.int32)
    if var:
        seq = torch.randint(ctx // 2, ctx + 1, (B,), device=DEV, generator=g, dtype=torch.int64)
    else:
        seq = torch.full((B,), ctx, dtype=torch.int64, device=DEV)
    rpi = torch.arange(B, dtype=torch.int64, device=DEV)
    loc = r2t.gather(1, (seq - 1)[:, None]).view(-1).to(torch.int64)
    kbufs = [torch.randn(unique + 1, 1, D, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
    q = torch.randn(B, Hq, D, device=DEV, dtype=torch.bfloat16, generator=g)
    k = torch.randn(B, 1, D, device=DEV, dtype=torch.bfloat16, generator=g)
    o = torch.empty(B, Hq, Dv, device=DEV, dtype=torch.bfloat16)
    S = ops.decode_attention_num_splits(B, Hq, 1, D, Dv, ctx, splits)
    ws = torch.empty(ops.decode_attention_workspace_bytes(B, Hq, Dv, S), dtype=torch.uint8, device=DEV)

    def run(i):
        ops.decode_attention(q, kbufs[i], kbufs[i].narrow(2, 0, Dv), o, k, k.narrow(2, 0, Dv), loc, ws, r2t, rpi, seq,
                             D ** -0.5, 0.0, max_seq_len=ctx, fused_kv_write=True, num_splits=splits)

    ms = timeit(run, layers, reps=10)
    tot = int(seq.sum())
    nbytes = tot * D * 2 + tot * 4 + B * Hq * (D + Dv) * 2 + B * D * 2 * 2
    report(f"mla {name} B{B} H{Hq} ctx{ctx} S{S}" + (" var" if var else "") + (" generic" if splits else ""), ms, nbytes,
           2 * tot * Hq * (D + Dv))
    del kbufs
    torch.cuda.empty_cache()


def bench_mla():
    mla_case("cfg5tp8", 256, 16, 16384)
    mla_case("cfg5tp8", 256, 16, 16384, var=True)
    mla_case("cfg5tp8-small", 64, 16, 4096, layers=4)
    mla_case("tp4", 64, 32, 8192)
    mla_case("tp1", 16, 128, 8192)
    mla_case("cfg5tp8-small", 64, 16, 4096, layers=4, splits=8)


def _plain_timeit():
    """no graph, 4 launches: for runs under ncu"""
    global timeit

    def timeit(fn, n_variants, reps=0, warm=3):  # noqa: F811
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(0); fn(1 % n_variants)
        e0.record()
        fn(0); fn(1 % n_variants)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 2


def bench_profmla():
    """a few plain launches of the cfg 5 per-GPU MLA shape for ncu (-k regex:decode_mla)"""
    _plain_timeit()
    mla_case("cfg5tp8", 256, 16, 16384)


def bench_profawq():
    """a few plain launches of the cfg 3 gate_up AWQ shape for ncu (-k regex:awq_fused)"""
    _plain_timeit()
    awq_case(128, 27648, 5120)


def main():
    what = sys.argv[1:] or ["attn", "gemm", "gemm_big", "int8", "awq"]
    print(f"# peaks: HBM {HBM} GB/s, bf16 {BF16_TF} TFLOP/s (MEASURED_PEAKS.json or fallback)")
    for w in what:
        print(f"## {w}")
        globals()[f"bench_{w}"]()


if __name__ == "__main__":
    main()
