"""TEST INFRASTRUCTURE ONLY -- loader for the reference's own CPU kernels (oracle/_ref).

`available()` is True when oracle/_ref/libsgl_ref_cpu.so exists (built by oracle/build_ref.py in the
authoring container, shipped prebuilt to the GPU box) AND the host CPU has the ISA it was compiled
for.  The reference's CPU GEMMs have no scalar path (gemm_fp8.cpp:100,120; gemm_int8.cpp:21), so on a
host without AVX512-BF16 + AMX the library is not loaded at all and callers fall back to the plain-C
port (oracle.cpu_oracle), saying so.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "_ref", "libsgl_ref_cpu.so")
REQUIRED_FLAGS = ("avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512_vnni", "avx512_bf16", "avx512_fp16",
                  "amx_tile", "amx_bf16", "amx_int8")
_loaded: Optional[bool] = None
_why = ""


def cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def available() -> bool:
    global _loaded, _why
    if _loaded is not None:
        return _loaded
    if not os.path.exists(SO_PATH):
        try:  # authoring container: build it from the reference tree if that is present
            from . import build_ref
            build_ref.build(verbose=False)
        except Exception as e:  # noqa: BLE001
            _why = f"build failed: {e}"
    if not os.path.exists(SO_PATH):
        _loaded, _why = False, _why or "oracle/_ref/libsgl_ref_cpu.so not built"
        return False
    missing = [f for f in REQUIRED_FLAGS if f not in cpu_flags()]
    if missing:
        _loaded, _why = False, f"host CPU lacks {','.join(missing)}"
        return False
    try:
        torch.ops.load_library(SO_PATH)
        _loaded = True
    except Exception as e:  # noqa: BLE001
        _loaded, _why = False, f"load failed: {e}"
    return _loaded


def why_unavailable() -> str:
    available()
    return _why


def ops():
    if not available():
        raise RuntimeError(f"reference CPU kernels unavailable: {_why}")
    return torch.ops.sgl_ref


def decode_attention_cpu(*args):
    """torch.ops.sgl_kernel.decode_attention_cpu of the reference (decode.cpp:1375)."""
    return ops().decode_attention_cpu(*args)


def extend_attention_cpu(*args):
    return ops().extend_attention_cpu(*args)
