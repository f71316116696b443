#!/usr/bin/env python
"""Developer tool: skinny vs large-M FP8 kernel at a given M (run once per B200_SK_MAX_M value).
    B200_SK_MAX_M=64 python tools/kb_cross.py 128"""
import sys

sys.argv, ms = sys.argv[:1], [int(a) for a in sys.argv[1:]] or [128]
import kbench  # noqa: E402

for m in ms:
    for n, k in [(3072, 4096), (4096, 2048), (14336, 4096), (4096, 7168), (1536, 4096), (4096, 1024), (7168, 4096), (4096, 3584)]:
        kbench.fp8_case(m, n, k)
