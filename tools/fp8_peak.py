#!/usr/bin/env python
"""Dense FP8 peak of this B200, measured the MEASURED_PEAKS.json way: torch._scaled_mm e4m3 x e4m3 ->
bf16 at 8192^3 (cuBLASLt, per-tensor scales), best of 10 (burst) and back to back for ~3 s (sustained), with the
clocks seen.  Also times our block-scaled kernel and, where it runs, vLLM's cutlass_scaled_mm with the same 1x128 /
128x128 block scales (the stand-in for the reference's sm100 CUTLASS kernel, fp8_blockwise_gemm_kernel.cu:165-343).

    python tools/fp8_peak.py > gpurun_out/fp8_peak.json
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

DEV = torch.device("cuda", 0)


def clocks():
    try:
        o = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader",
                            "-i", "0"], capture_output=True, text=True, timeout=10).stdout.strip()
        return o
    except Exception as e:  # noqa: BLE001
        return str(e)


def time_fn(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def burst_and_sustained(fn, flops, seconds=3.0):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = min(time_fn(fn, 1) for _ in range(10))
    n = max(10, int(seconds * 1e3 / best))
    t0 = time.time()
    sus = time_fn(fn, n)
    clk = clocks()
    return {"burst_tflops": round(flops / best / 1e9, 1), "sustained_tflops": round(flops / sus / 1e9, 1),
            "burst_ms": round(best, 4), "sustained_ms": round(sus, 4), "sustained_wall_s": round(time.time() - t0, 2),
            "clocks_after": clk}


def main():
    from iaas_sglang_b200 import ops

    out = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__}
    M = N = K = 8192
    flops = 2.0 * M * N * K
    g = torch.Generator(device=DEV).manual_seed(0)
    a = (torch.randn(M, K, device=DEV, generator=g)).to(torch.float8_e4m3fn)
    w = (torch.randn(N, K, device=DEV, generator=g)).to(torch.float8_e4m3fn)
    one = torch.ones((), device=DEV)
    try:
        out["cublaslt_fp8_per_tensor_8192"] = burst_and_sustained(
            lambda: torch._scaled_mm(a, w.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16), flops)
    except Exception as e:  # noqa: BLE001
        out["cublaslt_fp8_per_tensor_8192"] = {"error": str(e)[:300]}
    ab, wb = a.to(torch.bfloat16), w.to(torch.bfloat16)
    out["cublas_bf16_8192"] = burst_and_sustained(lambda: torch.matmul(ab, wb.t()), flops)
    del ab, wb
    sa = (torch.rand(K // 128, M, device=DEV, generator=g) * 1e-3 + 1e-4).t()
    sb = (torch.rand(N // 128, K // 128, device=DEV, generator=g) * 1e-3 + 1e-4).t()
    out["ours_fp8_blockwise_8192"] = burst_and_sustained(lambda: ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16), flops)
    # the kernel-to-beat stand-in: vLLM's CUTLASS block-scaled FP8 GEMM
    try:
        from vllm import _custom_ops as vops

        sa_c = sa.contiguous().t().contiguous().t() if False else sa   # [M, K/128] with stride(0) == 1 (M-major), as ours
        sb_v = sb                                                        # [K/128, N/128]
        fn = lambda: vops.cutlass_scaled_mm(a, w.t(), sa_c, sb_v, torch.bfloat16)  # noqa: E731
        ref = fn()
        mine = ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
        rel = ((ref.float() - mine.float()).norm() / mine.float().norm()).item()
        r = burst_and_sustained(fn, flops)
        r["rel_l2_vs_ours"] = rel
        out["vllm_cutlass_scaled_mm_blockwise_8192"] = r
    except Exception as e:  # noqa: BLE001
        out["vllm_cutlass_scaled_mm_blockwise_8192"] = {"error": str(e)[:400]}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
