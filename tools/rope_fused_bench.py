#!/usr/bin/env python
"""Developer tool: qkv GEMM + RoPE, fused epilogue against GEMM followed by the rope kernel (CUDA-graph timing)."""
import os
import sys

sys.argv = sys.argv[:1]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import kbench  # noqa: E402
import torch  # noqa: E402

from iaas_sglang_b200 import ops  # noqa: E402

for M, Hq, Hkv, K in ((64, 32, 8, 4096), (128, 16, 4, 4096), (32, 32, 8, 4096), (64, 8, 1, 8192)):
    D = 128
    N = (Hq + 2 * Hkv) * D
    g = torch.Generator(device=kbench.DEV).manual_seed(0)
    nv = max(1, min(16, kbench.L2_BYTES // (N * K) + 1))
    ws_ = [(torch.randn(N, K, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn) for _ in range(nv)]
    sb = (torch.rand((N + 127) // 128, (K + 127) // 128, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()
    a = (torch.randn(M, K, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sa = (torch.rand((K + 127) // 128, M, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()
    pos = torch.randint(0, 100000, (M,), device=kbench.DEV, dtype=torch.int64)
    cs = ops.rope_cos_sin(pos, D, 500000.0)
    rc = (Hq + Hkv) * D
    plain = kbench.timeit(lambda i: ops.fp8_blockwise_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16), nv) * 1e3
    two = kbench.timeit(lambda i: ops.fp8_blockwise_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16, rope=(pos, rc, D, 500000.0, None)), nv) * 1e3
    fused = kbench.timeit(lambda i: ops.fp8_blockwise_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16, rope=(pos, rc, D, 500000.0, cs)), nv) * 1e3
    tab = kbench.timeit(lambda i: ops.rope_cos_sin(pos, D, 500000.0, out=cs), 1) * 1e3
    print(f"M{M} N{N} K{K}: GEMM {plain:5.1f} us   GEMM + rope kernel {two:5.1f} us   fused {fused:5.1f} us   (cos/sin table, once per step: {tab:4.1f} us)", flush=True)
