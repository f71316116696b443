#!/usr/bin/env python
"""Developer tool: large-M FP8 block-scaled GEMM variants (knobs gb_pair / gb_bn) and probes (knob gb_epi: 1 no TMA,
2 no tcgen05.ld, 4 no FFMA, 8 no per-k-block promotion -- results are wrong with a probe bit set, only the time means
something).    python tools/gb_sweep.py [M N K]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from iaas_sglang_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda", 0)
M, N, K = [int(x) for x in sys.argv[1:4]] if len(sys.argv) >= 4 else (8192, 8192, 8192)
g = torch.Generator(device=dev).manual_seed(0)
w = (torch.randn(N, K, device=dev, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
a = (torch.randn(M, K, device=dev, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
sb = (torch.rand(N // 128, K // 128, device=dev, generator=g) * 1e-3 + 1e-4).t()
sa = (torch.rand(K // 128, M, device=dev, generator=g) * 1e-3 + 1e-4).t()


def run():
    return ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)


def timed(reps=10):
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


_lib.tune_set("sk_max_m", 0)
ref = None
if os.environ.get("ZERO"):
    a.view(torch.uint8).zero_(); w.view(torch.uint8).zero_()
for pair in (2, 4):
    for bn in (128, 256, 0):
        for dbg in (0,):
            _lib.tune_set("gb_pair", pair)
            _lib.tune_set("gb_bn", bn)
            _lib.tune_set("gb_epi", dbg)
            try:
                ms = timed()
                out = run()
                torch.cuda.synchronize()
                note = ""
                if dbg == 0:
                    if ref is None:
                        ref = out.float()
                    else:
                        note = f" rel_l2 vs first variant {((out.float() - ref).norm() / ref.norm()).item():.2e}"
                print(f"pair={'2cta' if pair == 4 else '1cta'} bn={bn} probe={dbg}: {ms * 1e3:8.1f} us  {2.0 * M * N * K / ms / 1e9:7.1f} TFLOP/s{note}", flush=True)
            except Exception as e:  # noqa: BLE001
                print(f"pair={pair} bn={bn} probe={dbg}: FAILED {e}", flush=True)
_lib.tune_set("gb_epi", 0)
# INT8 at the same shape (same kernel template, int32 accumulation in TMEM over the whole K loop)
ai = torch.randint(-127, 128, (M, K), device=dev, dtype=torch.int8)
wi = torch.randint(-127, 128, (N, K), device=dev, dtype=torch.int8)
si, sj = torch.rand(M, device=dev) * 1e-3, torch.rand(N, device=dev) * 1e-3
for pair, dbg in ((1, 0), (2, 0)):
    _lib.tune_set("gb_pair", pair)
    _lib.tune_set("gb_epi", dbg)
    run = lambda: ops.int8_scaled_mm(ai, wi.t(), si, sj, torch.bfloat16, None)  # noqa: E731
    ms = timed()
    print(f"int8 pair={pair} probe={dbg}: {ms * 1e3:8.1f} us  {2.0 * M * N * K / ms / 1e9:7.1f} TOP/s", flush=True)
_lib.tune_set("gb_epi", 0)
