"""Verbose GPU sanity sweep (prints error magnitudes); run on the GPU box: python tests/gpu_sanity.py"""
import os
import sys
import time
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from oracle import cpu_oracle as orc
from tests.util import DecodeCase, make_fp8_gemm_case, max_abs, rel_l2, run_cuda_decode


def section(name):
    print(f"\n=== {name} ===", flush=True)


def guard(fn, *a, **k):
    try:
        fn(*a, **k)
    except Exception:  # noqa: BLE001
        traceback.print_exc()
        print("FAILED", fn.__name__, a, k, flush=True)


def attn(B, Hq, Hkv, D, Dv, sl, **kw):
    fused = kw.pop("fused", False)
    cap = kw.pop("cap", 0.0)
    ns = kw.pop("num_splits", 0)
    c = DecodeCase(B, Hq, Hkv, D, Dv, sl, **kw)
    ref, kb, vb = c.oracle(logit_cap=cap)
    t = time.time()
    out, kb2, vb2 = run_cuda_decode(c, logit_cap=cap, fused=fused, num_splits=ns)
    print(f"attn B{B} Hq{Hq} Hkv{Hkv} D{D} Dv{Dv} sl={sl if isinstance(sl, int) else 'var'} fused={fused} cap={cap} ns={ns} {kw}: "
          f"relL2={rel_l2(out, ref):.2e} maxabs={max_abs(out, ref):.2e} kv_exact={torch.equal(kb, kb2) and torch.equal(vb, vb2)} "
          f"nan={bool(torch.isnan(out.float()).any())} ({time.time() - t:.2f}s)", flush=True)


def gemm_fp8(M, N, K):
    from iaas_sglang_b200 import ops
    a, b, sa, sb = make_fp8_gemm_case(M, N, K)
    ref = orc.fp8_blockwise_scaled_mm(a, b, sa, sb, torch.bfloat16)
    ad = a.cuda()
    bd = b.t().contiguous().cuda().t()
    sad = sa.t().contiguous().cuda().t()
    sbd = sb.t().contiguous().cuda().t()
    out = ops.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)
    torch.cuda.synchronize()
    print(f"fp8 gemm M{M} N{N} K{K}: relL2={rel_l2(out, ref):.2e} maxabs={max_abs(out, ref):.2e} refmax={ref.float().abs().max():.2e}", flush=True)


def gemm_int8(M, N, K, bias=True):
    from iaas_sglang_b200 import ops
    g = torch.Generator().manual_seed(0)
    a = (torch.randn(M, K, generator=g) * 5).clamp(-128, 127).round().to(torch.int8)
    b = (torch.randn(N, K, generator=g) * 5).clamp(-128, 127).round().to(torch.int8).t()
    sa, sb = torch.randn(M, generator=g), torch.randn(N, generator=g)
    bi = (torch.randn(N, generator=g) * 10).to(torch.bfloat16) if bias else None
    ref = orc.int8_scaled_mm(a, b, sa, sb, torch.bfloat16, bi)
    out = ops.int8_scaled_mm(a.cuda(), b.t().contiguous().cuda().t(), sa.cuda(), sb.cuda(), torch.bfloat16,
                             bi.cuda() if bias else None)
    torch.cuda.synchronize()
    print(f"int8 gemm M{M} N{N} K{K} bias={bias}: exact={torch.equal(out.cpu(), ref)} relL2={rel_l2(out, ref):.2e} maxabs={max_abs(out, ref):.2e}", flush=True)


def quant():
    from iaas_sglang_b200 import ops
    x = (torch.randn(64, 4096) * 3).to(torch.bfloat16)
    q, s = orc.per_token_group_quant_fp8(x)
    q2, s2 = ops.per_token_group_quant_fp8(x.cuda(), 128, column_major_scales=True)
    print("group quant fp8: q exact", torch.equal(q.view(torch.uint8), q2.cpu().view(torch.uint8)), "s exact", torch.equal(s, s2.cpu()),
          "s stride", s2.stride(), flush=True)
    q, s = orc.per_token_quant_int8(x)
    q2, s2 = ops.per_token_quant_int8(x.cuda())
    print("token quant int8: q exact", torch.equal(q, q2.cpu()), "s exact", torch.equal(s, s2.cpu()), flush=True)


def awq(K, N, M=16, dtype=torch.bfloat16):
    from iaas_sglang_b200 import ops
    g = torch.Generator().manual_seed(0)
    qw = torch.randint(0, 2**31 - 1, (K, N // 8), generator=g, dtype=torch.int32)
    qz = torch.randint(0, 2**31 - 1, (K // 128, N // 8), generator=g, dtype=torch.int32)
    sc = (torch.rand(K // 128, N, generator=g) * 0.01).to(dtype)
    ref = orc.awq_dequantize(qw, sc, qz)
    out = ops.awq_dequantize(qw.cuda(), sc.cuda(), qz.cuda())
    print(f"awq dequant K{K} N{N} {dtype}: exact={torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16))}", flush=True)
    x = (torch.randn(M, K, generator=g)).to(dtype)
    ref2 = orc.awq_gemm(x, qw, sc, qz)
    out2 = ops.awq_w4a16_gemm(x.cuda(), qw.cuda(), sc.cuda(), qz.cuda())
    torch.cuda.synchronize()
    print(f"awq w4a16 gemm M{M} K{K} N{N}: relL2={rel_l2(out2, ref2):.2e} maxabs={max_abs(out2, ref2):.2e}", flush=True)


def main():
    print(torch.cuda.get_device_name(0), torch.cuda.get_device_capability(0))
    os.system("lscpu | egrep 'Model name|^CPU\\(s\\)|Socket|Core|Thread' ; grep -o -E 'avx512_bf16|amx_bf16|amx_int8|avx512_fp16' /proc/cpuinfo | sort | uniq -c")
    from oracle import ref_ops
    print("oracle/_ref usable on this host:", ref_ops.available(), ref_ops.why_unavailable())
    section("quant")
    guard(quant)
    section("decode attention (tuned mma path)")
    guard(attn, 2, 32, 8, 128, 128, 300)
    guard(attn, 2, 32, 8, 128, 128, 300, fused=True)
    guard(attn, 16, 32, 8, 64, 64, 1024)
    guard(attn, 4, 8, 1, 128, 128, 1000, fused=True)
    guard(attn, 3, 20, 4, 128, 128, [1, 17, 555], fused=True)
    guard(attn, 3, 10, 2, 128, 128, [1, 17, 555])
    guard(attn, 2, 16, 16, 64, 64, 1024)
    guard(attn, 2, 64, 2, 128, 128, 200)          # G = 32 -> two q blocks
    guard(attn, 2, 32, 8, 128, 128, 2048, cap=30.0)
    guard(attn, 2, 32, 8, 128, 128, 700, dtype=torch.float16)
    guard(attn, 2, 32, 8, 128, 128, 700, idx_dtype=torch.int64, extra_reqs=5)
    guard(attn, 4, 32, 8, 128, 128, 600, shared_prefix=256, fused=True)
    guard(attn, 2, 32, 8, 128, 128, 513, noncontig=True)
    guard(attn, 2, 32, 8, 128, 128, 4096, num_splits=1)
    section("decode attention (generic path)")
    guard(attn, 2, 32, 8, 33, 55, 1024)
    guard(attn, 2, 16, 1, 16, 16, 1024)
    guard(attn, 2, 128, 1, 80, 80, 300)
    guard(attn, 1, 16, 1, 576, 512, 1024)
    guard(attn, 2, 16, 1, 576, 512, [5, 700], fused=True)
    guard(attn, 1, 22, 1, 576, 512, 100)
    guard(attn, 2, 128, 2, 512, 512, 100)
    section("fp8 blockwise gemm")
    for mnk in [(128, 128, 128), (128, 128, 512), (64, 256, 1024), (1, 128, 512), (5, 512, 1024), (127, 1024, 4096),
                (64, 4096, 4096), (512, 512, 512), (300, 384, 640)]:
        guard(gemm_fp8, *mnk)
    section("int8 gemm")
    for mnk in [(128, 128, 128), (16, 128, 512), (64, 4096, 4096), (1, 16, 512), (300, 136, 1024)]:
        guard(gemm_int8, *mnk)
    guard(gemm_int8, 32, 512, 1024, bias=False)
    section("awq")
    guard(awq, 512, 1024)
    guard(awq, 1024, 512, 64, torch.float16)
    print("\nsanity sweep done", flush=True)


if __name__ == "__main__":
    main()
