#!/usr/bin/env python
"""Developer tool: FP8 skinny (decode) kernel vs the persistent large-M kernel over the TP-local shapes of SURVEY 8a
row a11, in ONE process (knob sk_max_m through b200_tune_set).  Prints a table and the winner per (shape, M).

    python tools/kb_cross.py [M ...]        # default 32 64 96 128 160 192 256
"""
import json
import os
import sys

INT8 = "--int8" in sys.argv
ms_list = [int(a) for a in sys.argv[1:] if a != "--int8"] or [32, 64, 96, 128, 160, 192, 256]
sys.argv = sys.argv[:1]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import kbench  # noqa: E402
import torch  # noqa: E402

from iaas_sglang_b200 import _lib, ops  # noqa: E402

SHAPES = {
    "cfg2 tp1": [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)],
    "cfg2 tp4": [(1536, 4096), (4096, 1024), (7168, 4096), (4096, 3584)],
    "cfg4 tp8": [(1280, 8192), (8192, 1024), (7168, 8192), (8192, 3584)],
    "cfg5 tp8": [(2112, 7168), (3072, 1536), (7168, 2048), (4608, 7168), (7168, 2304)],
}


def case_int8(M, N, K):
    g = torch.Generator(device=kbench.DEV).manual_seed(0)
    nv = max(1, min(16, kbench.L2_BYTES // (N * K) + 1))
    ws_ = [torch.randint(-127, 128, (N, K), device=kbench.DEV, generator=g, dtype=torch.int8) for _ in range(nv)]
    a = torch.randint(-127, 128, (M, K), device=kbench.DEV, generator=g, dtype=torch.int8)
    sa = torch.rand(M, device=kbench.DEV, generator=g) * 1e-3
    sb = torch.rand(N, device=kbench.DEV, generator=g) * 1e-3

    def run(i):
        ops.int8_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16, None)

    out = {}
    for name, max_m in (("skinny", 256), ("large", 0), ("policy", -1)):
        _lib.tune_set("sk_max_m", max_m)
        out[name] = kbench.timeit(run, nv) * 1e3
    return out


def case(M, N, K):
    g = torch.Generator(device=kbench.DEV).manual_seed(0)
    nv = max(1, min(16, kbench.L2_BYTES // (N * K) + 1))
    ws_ = [(torch.randn(N, K, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn) for _ in range(nv)]
    sb = (torch.rand((N + 127) // 128, (K + 127) // 128, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()
    a = (torch.randn(M, K, device=kbench.DEV, generator=g) * 50).clamp_(-448, 448).to(torch.float8_e4m3fn)
    sa = (torch.rand((K + 127) // 128, M, device=kbench.DEV, generator=g) * 1e-3 + 1e-4).t()

    def run(i):
        ops.fp8_blockwise_scaled_mm(a, ws_[i].t(), sa, sb, torch.bfloat16)

    out = {}
    for name, max_m in (("skinny", 256), ("large", 0), ("policy", -1)):
        _lib.tune_set("sk_max_m", max_m)
        out[name] = kbench.timeit(run, nv) * 1e3
    return out


def main():
    old = _lib.tune_get("sk_max_m")
    rows = []
    for grp, shapes in SHAPES.items():
        for n, k in shapes:
            for m in ms_list:
                r = case_int8(m, n, k) if INT8 else case(m, n, k)
                floor = (n * k + m * k + m * n * 2) / kbench.HBM / 1e3
                win = "skinny" if r["skinny"] <= r["large"] else "large"
                best = min(r["skinny"], r["large"])
                flag = "  <-- policy off by %.0f%%" % (100 * (r["policy"] / best - 1)) if r["policy"] > 1.06 * best else ""
                print(f"{grp:9s} N{n:6d} K{k:6d} M{m:4d}  skinny {r['skinny']:7.1f} us  large {r['large']:7.1f} us  "
                      f"policy {r['policy']:7.1f} us  hbm floor {floor:6.1f} us  -> {win}{flag}", flush=True)
                rows.append(dict(group=grp, N=n, K=k, M=m, skinny_us=round(r["skinny"], 2), large_us=round(r["large"], 2),
                                 hbm_floor_us=round(floor, 2)))
    _lib.tune_set("sk_max_m", old)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/kb_cross.json", "w") as f:
        json.dump(rows, f)


if __name__ == "__main__":
    main()
