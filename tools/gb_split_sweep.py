#!/usr/bin/env python
"""Developer tool: large-M fp8 kernel with and without split-K (knob gb_split) on the few-tile TP-local shapes."""
import os
import sys

sys.argv = sys.argv[:1]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import kbench  # noqa: E402
import torch  # noqa: E402

from iaas_sglang_b200 import _lib, ops  # noqa: E402

CASES = [("cfg2 tp4", 256, [(1536, 4096), (4096, 1024), (7168, 4096), (4096, 3584)]),
         ("cfg2 tp8", 512, [(768, 4096), (4096, 512), (3584, 4096), (4096, 1792)]),
         ("cfg4 tp8", 256, [(1280, 8192), (8192, 1024), (7168, 8192), (8192, 3584)]),
         ("cfg4 tp4", 256, [(2560, 8192), (8192, 2048), (14336, 8192), (8192, 7168)]),
         ("cfg5 tp8", 256, [(2112, 7168), (3072, 1536), (7168, 2048), (4608, 7168), (7168, 2304)]),
         ("cfg2 tp2", 128, [(4096, 2048), (3072, 4096)])]


def main():
    for grp, M, shapes in CASES:
        for n, k in shapes:
            g = torch.Generator(device=kbench.DEV).manual_seed(0)
            nv = max(1, min(16, kbench.L2_BYTES // (n * k) + 1))
            ws_ = [(torch.randn(n, k, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn) for _ in range(nv)]
            sb = (torch.rand((n + 127) // 128, (k + 127) // 128, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()
            a = (torch.randn(M, k, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
            sa = (torch.rand((k + 127) // 128, M, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()

            def run(i):
                return ops.fp8_blockwise_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16)

            _lib.tune_set("sk_max_m", 0)     # large-M kernel
            line = f"{grp} M{M:4d} N{n:6d} K{k:6d} |"
            ref = None
            for sp in (1, 0, 2, 4, 8):
                _lib.tune_set("gb_split", sp)
                out = run(0)
                if ref is None:
                    ref = out.float()
                err = ((out.float() - ref).norm() / ref.norm()).item()
                line += f" split{sp} {kbench.timeit(run, nv) * 1e3:6.1f}us (rel {err:.1e})"
            _lib.tune_set("gb_split", 0)
            _lib.tune_set("sk_max_m", -1)
            line += f" | policy {kbench.timeit(run, nv) * 1e3:6.1f}us"
            print(line, flush=True)


if __name__ == "__main__":
    main()
