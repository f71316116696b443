"""Build iaas_sglang_b200/csrc/libb200sgl.so (the C-ABI library) with nvcc for sm_100a, in-tree.

    python iaas_sglang_b200/csrc/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels to the GPU box with gpurun.
"""
from __future__ import annotations

import os
import subprocess
import sys
import time
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(HERE, "libb200sgl.so")
SOURCES = ["runtime.cu", "elementwise.cu", "decode_attn.cu", "extend_attn.cu", "gemm_tc.cu", "gemm_skinny.cu", "gemm_big.cu", "awq_gemm.cu", "awq_fused.cu", "fused_elem.cu", "allreduce.cu", "cascade.cu"]
HEADERS = ["common.cuh", "gemm_skinny.h", "decode_mla.cuh", "rmsnorm_row.cuh", os.path.join(ROOT, "include", "b200_sgl.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr"]
FLAGS += os.environ.get("B200_NVCC_EXTRA", "").split()


def sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(HERE, s))]


def up_to_date() -> bool:
    if not os.path.exists(SO):
        return False
    t = os.path.getmtime(SO)
    deps = [os.path.join(HERE, s) for s in sources()] + [h if os.path.isabs(h) else os.path.join(HERE, h) for h in HEADERS]
    deps.append(os.path.abspath(__file__))
    return all(os.path.getmtime(d) <= t for d in deps)


def _compile(src: str, verbose: bool, force: bool = False):
    obj = os.path.join(HERE, "build", src.replace(".cu", ".o"))
    deps = [os.path.join(HERE, src), os.path.abspath(__file__)] + [h if os.path.isabs(h) else os.path.join(HERE, h) for h in HEADERS]
    if not force and not verbose and os.path.exists(obj) and all(os.path.getmtime(d) <= os.path.getmtime(obj) for d in deps):
        return src, obj, 0, ""   # incremental: this object is newer than its source and every header
    cmd = [NVCC, *FLAGS, "-c", os.path.join(HERE, src), "-o", obj]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return src, obj, r.returncode, r.stdout


def build(force: bool = False, verbose: bool = False) -> str:
    """Serialised across processes (torchrun starts N ranks that may all find the sources newer): an exclusive
    file lock around the whole build, objects and the link go to temporary names and are renamed into place, so a
    rank that is dlopen-ing the library never sees a half-written file."""
    import fcntl

    if not force and up_to_date():
        return SO
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}; cannot build {SO}")
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    with open(os.path.join(HERE, "build", ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and up_to_date():   # another process built it while we waited
                return SO
            return _build_locked(verbose, force)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool, force: bool = False) -> str:
    t0 = time.time()
    with ThreadPoolExecutor(max_workers=8) as ex:
        results = list(ex.map(lambda s: _compile(s, verbose, force), sources()))
    objs = []
    for src, obj, rc, out in results:
        if verbose or rc != 0:
            sys.stderr.write(f"--- {src} ---\n{out}\n")
        if rc != 0:
            raise RuntimeError(f"nvcc failed on {src}")
        objs.append(obj)
    tmp = f"{SO}.{os.getpid()}.tmp"
    cmd = [NVCC, "-shared", "-o", tmp, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "shared"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}")
    os.replace(tmp, SO)
    if verbose:
        sys.stderr.write(f"built {SO} in {time.time() - t0:.1f}s\n")
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
