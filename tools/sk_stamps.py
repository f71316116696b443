"""developer: per-phase time stamps of the skinny GEMM (B200_SK_DEBUG=256), no CUDA graph"""
import os, sys
os.environ["B200_SK_DEBUG"] = os.environ.get("B200_SK_DEBUG", "256")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iaas_sglang_b200 import ops
dev = torch.device("cuda", 0)
for (M, N, K) in [(64, 4096, 4096), (256, 8192, 1024), (256, 7168, 8192), (256, 1280, 8192), (128, 7168, 8192)]:
    w = (torch.randn(N, K, device=dev) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    a = (torch.randn(M, K, device=dev) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sb = torch.rand(N // 128, K // 128, device=dev) * 1e-3
    sa = (torch.rand(K // 128, M, device=dev) * 1e-3).t()
    for _ in range(8):
        ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb.t(), torch.bfloat16)
    torch.cuda.synchronize()
