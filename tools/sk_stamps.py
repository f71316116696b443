#!/usr/bin/env python
"""Developer tool: time stamps inside the decode-shape FP8 GEMM (knob sk_debug = 256; gemm_skinny.cu SK_STAMP):
0 CTA start, 1 epilogue warps ready, 2 k-loop done (all MMAs seen), 3 cluster barrier A passed, 4 DSMEM copies issued,
5 partials received, 6 stores done."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from iaas_sglang_b200 import _lib, ops  # noqa: E402

M = int(os.environ.get("SK_M", "64"))
for n, k, s in ((4096, 4096, 0), (4096, 4096, 1), (4096, 14336, 0), (4096, 14336, 8), (28672, 4096, 1)):
    g = torch.Generator(device="cuda").manual_seed(0)
    w = (torch.randn(n, k, device="cuda", generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sb = (torch.rand((n + 127) // 128, (k + 127) // 128, device="cuda", generator=g) * 1e-3 + 1e-4).t()
    a = (torch.randn(M, k, device="cuda", generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sa = (torch.rand((k + 127) // 128, M, device="cuda", generator=g) * 1e-3 + 1e-4).t()
    _lib.tune_set("sk_max_m", 256)
    _lib.tune_set("sk_splits", s)
    _lib.tune_set("sk_debug", 256)
    for _ in range(8):
        ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
    torch.cuda.synchronize()
    _lib.tune_set("sk_debug", 0)
