#!/usr/bin/env python
"""Developer tool: decode-shape FP8 GEMMs (M 64) of cfg 2 -- skinny kernel by forced split count / ring depth, large-M
kernel by tile width -- CUDA-graph timing like kbench (knobs through b200_tune_set)."""
import os
import sys

sys.argv = sys.argv[:1]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import kbench  # noqa: E402
import torch  # noqa: E402

from iaas_sglang_b200 import _lib, ops  # noqa: E402

SHAPES = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]
M = int(os.environ.get("SK_M", "64"))


def main():
    for n, k in SHAPES:
        g = torch.Generator(device=kbench.DEV).manual_seed(0)
        nv = max(1, min(16, kbench.L2_BYTES // (n * k) + 1))
        ws_ = [(torch.randn(n, k, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn) for _ in range(nv)]
        sb = (torch.rand((n + 127) // 128, (k + 127) // 128, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()
        a = (torch.randn(M, k, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
        sa = (torch.rand((k + 127) // 128, M, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()

        def run(i):
            ops.fp8_blockwise_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16)

        floor = (n * k + M * k + M * n * 2) / kbench.HBM / 1e3
        line = f"N{n:6d} K{k:6d} M{M:4d} floor {floor:5.1f} |"
        _lib.tune_set("sk_max_m", 256)
        for s in (0, 1, 2, 3, 4, 6, 8):
            _lib.tune_set("sk_splits", s)
            try:
                line += f" S{s} {kbench.timeit(run, nv) * 1e3:5.1f}"
            except Exception as e:  # noqa: BLE001
                line += f" S{s}  err"
        _lib.tune_set("sk_splits", 0)
        for st in (4, 6, 8):
            _lib.tune_set("sk_stages", st)
            line += f" st{st} {kbench.timeit(run, nv) * 1e3:5.1f}"
        _lib.tune_set("sk_stages", 0)
        _lib.tune_set("sk_max_m", 0)
        for bn in (128, 256):
            _lib.tune_set("gb_bn", bn)
            line += f" | big{bn} {kbench.timeit(run, nv) * 1e3:5.1f}"
        _lib.tune_set("gb_bn", 0)
        _lib.tune_set("sk_max_m", -1)
        line += f" | policy {kbench.timeit(run, nv) * 1e3:5.1f}"
        print(line, flush=True)


if __name__ == "__main__":
    main()
