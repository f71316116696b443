import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import torch
from iaas_sglang_b200 import ops
DEV = torch.device("cuda", 0)
B, Hq, Hkv, ctx, D = 64, 32, 8, 4096, 128
g = torch.Generator(device=DEV).manual_seed(1)
T = B * ctx
r2t = (torch.randperm(T, device=DEV, generator=g) + 1).to(torch.int32).view(B, ctx)
seq = torch.full((B,), ctx, dtype=torch.int64, device=DEV); rpi = torch.arange(B, dtype=torch.int64, device=DEV)
q = torch.randn(B, Hq, D, device=DEV, dtype=torch.bfloat16, generator=g); o = torch.empty_like(q)
kb = [torch.randint(0, 120, (T + 1, Hkv, D), device=DEV, dtype=torch.uint8, generator=g) for _ in range(2)]
vb = [torch.randint(0, 120, (T + 1, Hkv, D), device=DEV, dtype=torch.uint8, generator=g) for _ in range(2)]
for i in range(4):
    ops.decode_attention(q, kb[i % 2], vb[i % 2], o, None, None, None, None, r2t, rpi, seq, D ** -0.5, 0.0, max_seq_len=ctx, k_scale=0.05, v_scale=0.05)
torch.cuda.synchronize()
