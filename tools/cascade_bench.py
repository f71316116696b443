"""developer: where the time of cascade decode goes (prefix kernel / suffix kernel + merge) vs plain decode attention"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iaas_sglang_b200 import ops

dev = torch.device("cuda", 0)


def ev(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def case(B, Hq, Hkv, D, ctx, shared):
    g = torch.Generator(device=dev).manual_seed(1)
    unique = B * (ctx - shared) + shared
    perm = (torch.randperm(unique, device=dev, generator=g) + 1).to(torch.int32)
    r2t = torch.zeros(B, ctx, dtype=torch.int32, device=dev)
    r2t[:, :shared] = perm[:shared][None]
    r2t[:, shared:] = perm[shared:].view(B, ctx - shared)
    kb = torch.randn(unique + 1, Hkv, D, device=dev, dtype=torch.bfloat16, generator=g)
    vb = torch.randn(unique + 1, Hkv, D, device=dev, dtype=torch.bfloat16, generator=g)
    q = torch.randn(B, Hq, D, device=dev, dtype=torch.bfloat16, generator=g)
    seq = torch.full((B,), ctx, dtype=torch.int64, device=dev)
    rpi = torch.arange(B, dtype=torch.int64, device=dev)
    o = torch.empty(B, Hq, D, device=dev, dtype=torch.bfloat16)
    st = ops.CascadeState(B, Hq, D, torch.bfloat16, dev)
    st.detect(r2t, rpi, seq)
    t_det = ev(lambda: st.detect(r2t, rpi, seq))
    t_plain = ev(lambda: ops.decode_attention(q, kb, vb, o, None, None, None, None, r2t, rpi, seq, D ** -0.5, 0.0, max_seq_len=ctx))
    ref = o.clone()
    t_pre = ev(lambda: st.prefix_attention(q, kb, vb, r2t, D ** -0.5, 0.0))
    t_suf = ev(lambda: ops.decode_attention(q, kb, vb, o, None, None, None, None, r2t, rpi, seq, D ** -0.5, 0.0, max_seq_len=ctx, cascade=st))
    rel = ((o.float() - ref.float()).norm() / ref.float().norm()).item()
    flops = 2.0 * B * Hq * int(st.prefix_len.item()) * D * 2
    print(f"B{B} Hq{Hq} Hkv{Hkv} D{D} ctx{ctx} shared{shared} (P={int(st.prefix_len.item())}): plain {t_plain:7.1f} us | detect {t_det:5.1f}  prefix {t_pre:7.1f} "
          f"({flops / t_pre / 1e6:6.1f} TFLOP/s)  suffix+merge {t_suf:7.1f}  = {t_pre + t_suf:7.1f} us | rel {rel:.1e}", flush=True)


case(256, 8, 1, 128, 8192, 2048)
case(256, 64, 8, 128, 8192, 2048)
case(256, 16, 2, 128, 8192, 2048)
case(64, 32, 8, 128, 4096, 2048)
case(256, 8, 1, 128, 8192, 6144)
