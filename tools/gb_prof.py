"""developer: pipeline-wait breakdown of the large-M fp8 GEMM (needs a -DGB_PROF probe build, tools/build_probe.sh:
B200_SGL_LIB=iaas_sglang_b200/csrc/libb200sgl_p0.so python tools/gb_prof.py), no CUDA graph"""
import os, sys
os.environ["B200_GB_PROF"] = "1"
os.environ["B200_SK_MAX_M"] = "0"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iaas_sglang_b200 import ops, _lib
dev = torch.device("cuda", 0)
M, N, K = 8192, 8192, 8192
w = (torch.randn(N, K, device=dev) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
a = (torch.randn(M, K, device=dev) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
sb = torch.rand(N // 128, K // 128, device=dev) * 1e-3
sa = (torch.rand(K // 128, M, device=dev) * 1e-3).t()
for bn in [int(x) for x in sys.argv[1:]] or [256]:
    _lib.tune_set("gb_bn", bn)
    best = 1e9
    for _ in range(12):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb.t(), torch.bfloat16)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"{os.environ.get('B200_SGL_LIB', 'default')} bn={bn}: (with the profiler's sync per launch) {2.0 * M * N * K / best / 1e9:.0f} TFLOP/s", file=sys.stderr)
