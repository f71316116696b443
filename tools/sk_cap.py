import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import torch
from iaas_sglang_b200 import _lib, ops
import kbench
M=64
for n,k in ((4096,14336),(6144,4096)):
    g = torch.Generator(device="cuda").manual_seed(0)
    w = (torch.randn(n, k, device="cuda", generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sb = (torch.rand((n + 127) // 128, (k + 127) // 128, device="cuda", generator=g) * 1e-3 + 1e-4).t()
    a = (torch.randn(M, k, device="cuda", generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sa = (torch.rand((k + 127) // 128, M, device="cuda", generator=g) * 1e-3 + 1e-4).t()
    _lib.tune_set("sk_max_m", 256)
    _lib.tune_set("sk_debug", 64)
    ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
    torch.cuda.synchronize()
    _lib.tune_set("sk_debug", 0)
    line=f"N{n} K{k}:"
    for s in (4,5,6,7,8):
        _lib.tune_set("sk_splits", s)
        line += f" S{s} {kbench.timeit(lambda i: ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16), 1)*1e3:5.1f}"
    _lib.tune_set("sk_splits", 0)
    print(line, flush=True)
