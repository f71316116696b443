#!/usr/bin/env python
"""Turn the ncu outputs brought back in gpurun_out/ into the tracked summaries under profiles/.

    python tools/summarize_profile.py launches gpurun_out/launches_r01.csv profiles/r01_launches.md [first_kernel_regex]
    python tools/summarize_profile.py kernel   gpurun_out/prof_x.ncu-rep   profiles/r01_x.md

`launches`: per-kernel share of one decode step from the `--metrics gpu__time_duration.sum` launch list
(cold-cache, serialised times: the SHARES are what is comparable with bench.py, not the absolutes).
`kernel`: the headline counters of an `ncu --set full` capture plus the top stall sites of the source page.
"""
from __future__ import annotations

import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict


def launches(src: str, dst: str, start_re: str = r"rmsnorm_kernel"):
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    data = rows[1:]
    ours = [r for r in data if "b200::" in r[ki] or re.match(r"(void )?(decode_|gemm_|rmsnorm|group_quant|rope_|silu_|awq_|per_token|set_kv)", r[ki])]
    # one step = from the first rmsnorm of a pass to the next pass's first rmsnorm; use the LAST complete pass
    names = [re.sub(r"\(.*", "", r[ki]).replace("void ", "") for r in ours]
    agg: "OrderedDict[str, list]" = OrderedDict()
    for n, r in zip(names, ours):
        agg.setdefault(n, []).append(float(r[vi].replace(",", "")))
    total = sum(sum(v) for v in agg.values())
    out = io.StringIO()
    out.write(f"# Launch list summary ({src})\n\n")
    out.write("`ncu --metrics gpu__time_duration.sum --clock-control none`; times are cold-cache and serialised, so the\n"
              "share of each kernel in the step is what to compare with bench.py, not the absolute.\n\n")
    out.write(f"{len(data)} launches captured, {len(ours)} from libb200sgl.so (the rest is torch setup: randn / fill / copy).\n\n")
    out.write("| kernel | launches | mean us | total us | share of our kernels |\n|---|---:|---:|---:|---:|\n")
    for n, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        out.write(f"| `{n}` | {len(v)} | {sum(v) / len(v) / 1e3:.1f} | {sum(v) / 1e3:.1f} | {sum(v) / total:.1%} |\n")
    out.write("\n## One layer, in launch order (first full layer found)\n\n| # | kernel | grid | block | us |\n|---|---|---|---|---:|\n")
    started = 0
    for r in data:
        if r not in ours:
            continue
        nm = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
        if re.search(start_re, r[ki]):
            started += 1
            if started == 3:
                break
        if started:
            out.write(f"| {r[0]} | `{nm}` | {r[gi]} | {r[bi]} | {float(r[vi].replace(',', '')) / 1e3:.1f} |\n")
    open(dst, "w").write(out.getvalue())
    print(out.getvalue())


KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
]


def kernel(src: str, dst: str, topn: int = 25):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = io.StringIO()
    out.write(f"# ncu --set full summary ({src})\n\n")
    ni = hdr.index("Kernel Name")
    for r in rows[2:]:
        out.write(f"## `{r[ni][:110]}`\n\n| metric | value | unit |\n|---|---:|---|\n")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                out.write(f"| {k} | {r[i]} | {units[i]} |\n")
        out.write("\n")
    srcp = subprocess.run(["ncu", "-i", src, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(srcp)))
    blocks, cur = [], None
    for r in srows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            blocks.append(cur)
        elif cur is not None:
            cur["rows"].append(r)
    for b in blocks[:1]:
        h = b["rows"][0]
        si, so = h.index("# Samples"), h.index("Source")
        data = [r for r in b["rows"][1:] if len(r) > si]
        tot = sum(int(r[si]) for r in data) or 1
        stall_cols = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
        agg = {}
        for r in data:
            for c in stall_cols:
                agg[h[c][6:]] = agg.get(h[c][6:], 0) + int(r[c])
        out.write(f"## Warp-state samples (source page, first captured launch): {tot} samples\n\n")
        out.write("by reason: " + ", ".join(f"{k} {v / tot:.1%}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]) + "\n\n")
        out.write("| SASS idx | samples | share | instruction | top reasons |\n|---:|---:|---:|---|---|\n")
        top = sorted(range(len(data)), key=lambda i: -int(data[i][si]))[:topn]
        for i in sorted(top):
            r = data[i]
            st = {h[c][6:]: int(r[c]) for c in stall_cols if int(r[c]) > 0}
            st = ", ".join(f"{k} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
            out.write(f"| {i} | {r[si]} | {int(r[si]) / tot:.1%} | `{r[so].strip()[:70]}` | {st} |\n")
    open(dst, "w").write(out.getvalue())
    print(out.getvalue())


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "launches":
        launches(*sys.argv[2:5])
    else:
        kernel(sys.argv[2], sys.argv[3])
