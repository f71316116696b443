"""TEST INFRASTRUCTURE ONLY -- build the reference's own CPU kernels into oracle/_ref/.

Compiles the reference CPU sources *where they lie* under /root/reference
(sgl-kernel/csrc/cpu/{decode,extend,gemm,gemm_fp8,gemm_int8}.cpp, unmodified, never copied)
plus oracle/ref_binder.cpp into oracle/_ref/libsgl_ref_cpu.so, a torch-ops library that
exposes torch.ops.sgl_ref.{decode_attention_cpu, extend_attention_cpu, fp8_scaled_mm_cpu, ...}.

The output directory is git-ignored but travels to the GPU box (it is not in .gpurunignore).
/root/reference does not exist on the GPU box: there only the prebuilt .so is used, and only if
the host CPU has the ISA it was compiled for (AVX512-BF16 + AMX), see oracle/ref_ops.py.

Recipe: g++ -O3 -fopenmp -std=c++17 with an explicit Sapphire-Rapids
ISA list (not -march=native, so the artefact does not depend on the build host), link -lgomp.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("B200_SGL_REFERENCE", "/root/reference")
CPU_DIR = os.path.join(REF_ROOT, "sgl-kernel", "csrc", "cpu")
OUT_DIR = os.path.join(HERE, "_ref")
OUT_SO = os.path.join(OUT_DIR, "libsgl_ref_cpu.so")
SOURCES = ["decode.cpp", "extend.cpp", "gemm.cpp", "gemm_fp8.cpp", "gemm_int8.cpp"]

ISA_FLAGS = [
    "-mavx512f", "-mavx512bw", "-mavx512vl", "-mavx512dq", "-mavx512cd", "-mavx512vnni",
    "-mavx512bf16", "-mavx512fp16", "-mamx-tile", "-mamx-bf16", "-mamx-int8", "-mfma", "-mf16c", "-mavx2",
]


def reference_available() -> bool:
    return all(os.path.exists(os.path.join(CPU_DIR, s)) for s in SOURCES)


def up_to_date() -> bool:
    if not os.path.exists(OUT_SO):
        return False
    if not reference_available():
        return True  # GPU box: use the prebuilt artefact as is
    t = os.path.getmtime(OUT_SO)
    deps = [os.path.join(CPU_DIR, s) for s in SOURCES] + [os.path.join(HERE, "ref_binder.cpp"), __file__]
    return all(os.path.getmtime(d) <= t for d in deps)


def build(verbose: bool = True, jobs: int = 6) -> str | None:
    if up_to_date():
        return OUT_SO
    if not reference_available():
        if verbose:
            print("[oracle/_ref] reference sources not present; nothing to build", file=sys.stderr)
        return None
    import torch
    from torch.utils import cpp_extension

    os.makedirs(OUT_DIR, exist_ok=True)
    inc = cpp_extension.include_paths() + [
        CPU_DIR,
        os.path.join(REF_ROOT, "sgl-kernel", "include"),
        sysconfig.get_paths()["include"],
    ]
    cflags = [
        "-O3", "-fopenmp", "-std=c++17", "-fPIC", "-Wno-unknown-pragmas", "-Wno-ignored-attributes",
        "-Wno-deprecated-declarations", "-DTORCH_API_INCLUDE_EXTENSION_H",
        f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
    ] + ISA_FLAGS
    for i in inc:
        cflags += ["-isystem", i]
    objs, procs = [], []
    srcs = [os.path.join(CPU_DIR, s) for s in SOURCES] + [os.path.join(HERE, "ref_binder.cpp")]
    t0 = time.time()
    for s in srcs:
        o = os.path.join(OUT_DIR, os.path.basename(s).replace(".cpp", ".o"))
        objs.append(o)
        procs.append((s, subprocess.Popen(["g++", *cflags, "-c", s, "-o", o],
                                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"[oracle/_ref] compile failed for {s}:\n{out.decode()[-4000:]}")
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    link = [
        "g++", "-shared", "-o", OUT_SO, *objs, f"-L{torch_lib}", "-ltorch", "-ltorch_cpu", "-lc10",
        "-L/usr/lib/gcc/x86_64-linux-gnu/13", "-lgomp", f"-Wl,-rpath,{torch_lib}",
    ]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError(f"[oracle/_ref] link failed:\n{r.stdout.decode()[-4000:]}")
    for o in objs:
        os.remove(o)
    if verbose:
        print(f"[oracle/_ref] built {OUT_SO} in {time.time() - t0:.1f}s", file=sys.stderr)
    return OUT_SO


if __name__ == "__main__":
    print(build())
