"""developer: the small kernels between the GEMMs (rmsnorm+quant, silu*mul+quant, rope, group quant) at the BASELINE shapes"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import kbench
import torch
from iaas_sglang_b200 import ops

dev = kbench.DEV
for M, H in ((64, 4096), (128, 4096), (512, 4096), (256, 8192), (128, 5120), (256, 7168)):
    nv = 8
    xs = [torch.randn(M, H, device=dev, dtype=torch.bfloat16) for _ in range(nv)]
    rs = [torch.randn(M, H, device=dev, dtype=torch.bfloat16) for _ in range(nv)]
    w = torch.ones(H, device=dev, dtype=torch.bfloat16)
    ms = kbench.timeit(lambda i: ops.rmsnorm_quant_fp8(xs[i], w, 1e-5, residual=rs[i]), nv)
    nbytes = M * H * (2 + 2 + 2 + 1) + M * H // 128 * 4
    kbench.report(f"rmsnorm_quant_fp8 +residual M{M} H{H}", ms, nbytes, 0)
    ms = kbench.timeit(lambda i: ops.rmsnorm(xs[i], w, 1e-5, residual=rs[i]), nv)
    kbench.report(f"rmsnorm +residual M{M} H{H}", ms, M * H * 8, 0)
for M, d in ((64, 14336), (512, 1792), (256, 3584)):
    xs = [torch.randn(M, 2 * d, device=dev, dtype=torch.bfloat16) for _ in range(8)]
    ms = kbench.timeit(lambda i: ops.silu_and_mul_quant_fp8(xs[i]), 8)
    kbench.report(f"silu_and_mul_quant_fp8 M{M} d{d}", ms, M * d * 5, 0)
