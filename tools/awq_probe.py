"""Developer probe: one fused AWQ GEMM call per process, so that a hung kernel costs one `timeout`.

    python tools/awq_probe.py M N K [group]      (env: B200_AWQ_NT, B200_AWQ_SPLITS, B200_AWQ_DEBUG=1)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from iaas_sglang_b200 import ops  # noqa: E402
from oracle import cpu_oracle as orc  # noqa: E402

M, N, K = (int(v) for v in sys.argv[1:4])
group = int(sys.argv[4]) if len(sys.argv) > 4 else 128
g = torch.Generator().manual_seed(M + N + K)
qw = torch.randint(-2 ** 31, 2 ** 31 - 1, (K, N // 8), generator=g, dtype=torch.int32)
qz = torch.randint(-2 ** 31, 2 ** 31 - 1, (K // group, N // 8), generator=g, dtype=torch.int32)
sc = (torch.rand(K // group, N, generator=g) * 0.01 + 1e-3).to(torch.bfloat16)
x = torch.randn(M, K, generator=g).to(torch.bfloat16)
ref = x.float() @ orc.awq_dequantize(qw, sc, qz).float()
out = ops.awq_w4a16_gemm(x.cuda(), qw.cuda(), sc.cuda(), qz.cuda())
torch.cuda.synchronize()
err = ((out.cpu().float() - ref).norm() / ref.norm()).item()
print(f"M={M} N={N} K={K} NT={os.environ.get('B200_AWQ_NT')} S={os.environ.get('B200_AWQ_SPLITS')} rel_l2={err:.3e}", flush=True)
