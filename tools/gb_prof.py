"""developer: pipeline-wait breakdown of the large-M fp8 GEMM (B200_GB_PROF=1), no CUDA graph"""
import os, sys
os.environ["B200_GB_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iaas_sglang_b200 import ops
dev = torch.device("cuda", 0)
M, N, K = 4096, 8192, 8192
w = (torch.randn(N, K, device=dev) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
a = (torch.randn(M, K, device=dev) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
sb = torch.rand(N // 128, K // 128, device=dev) * 1e-3
sa = (torch.rand(K // 128, M, device=dev) * 1e-3).t()
for _ in range(12):
    ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb.t(), torch.bfloat16)
torch.cuda.synchronize()
