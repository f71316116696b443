"""developer: decode attention over an FP8 (e4m3) KV pool vs the bf16 pool, same shapes (bytes halve)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import kbench
import torch
from iaas_sglang_b200 import ops

DEV = kbench.DEV


def case(name, B, Hq, Hkv, ctx, layers=3):
    D = 128
    g = torch.Generator(device=DEV).manual_seed(1)
    T = B * ctx
    r2t = (torch.randperm(T, device=DEV, generator=g) + 1).to(torch.int32).view(B, ctx)
    seq = torch.full((B,), ctx, dtype=torch.int64, device=DEV)
    rpi = torch.arange(B, dtype=torch.int64, device=DEV)
    q = torch.randn(B, Hq, D, device=DEV, dtype=torch.bfloat16, generator=g)
    o = torch.empty_like(q)
    for fp8 in (False, True):
        if fp8:
            kb = [torch.randint(0, 120, (T + 1, Hkv, D), device=DEV, dtype=torch.uint8, generator=g) for _ in range(layers)]
            vb = [torch.randint(0, 120, (T + 1, Hkv, D), device=DEV, dtype=torch.uint8, generator=g) for _ in range(layers)]
            kw = dict(k_scale=0.05, v_scale=0.05)
        else:
            kb = [torch.randn(T + 1, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
            vb = [torch.randn(T + 1, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
            kw = {}
        ms = kbench.timeit(lambda i: ops.decode_attention(q, kb[i], vb[i], o, None, None, None, None, r2t, rpi, seq, D ** -0.5, 0.0,
                                                          max_seq_len=ctx, **kw), layers, reps=12)
        eb = 1 if fp8 else 2
        nbytes = T * Hkv * 2 * D * eb + T * 4 + B * Hq * 2 * D * 2
        kbench.report(f"attn {name} {'e4m3' if fp8 else 'bf16'} KV B{B} H{Hq}/{Hkv} ctx{ctx}", ms, nbytes, 2 * T * Hq * 2 * D)
        del kb, vb
        torch.cuda.empty_cache()


case("cfg2", 64, 32, 8, 4096)
case("cfg3tp2", 128, 20, 4, 4096)
case("cfg2tp8", 512, 4, 1, 4096)
case("cfg4tp8", 256, 8, 1, 8192)
