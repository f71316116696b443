"""GPU tier, needs >= 2 visible GPUs (skipped otherwise): the custom all-reduce kernels that carry every N > 1 number
(csrc/allreduce.cu) -- rank-order fp32 sum, identical on all ranks, mixed sizes with a delayed rank, CUDA-graph replay,
and the fused all-reduce + residual-add + RMSNorm (+ FP8 quant) against the two-kernel sequence.  The checks live in
tools/ar_check.py (one process per GPU under torch.distributed.run); this test spawns it on 2 ranks (and on every GPU
of the box when there are more than 2)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(n):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "ar_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    tail = "\n".join(r.stdout.splitlines()[-40:])
    assert r.returncode == 0 and "AR_CHECK PASS" in r.stdout, tail


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_allreduce_kernels_two_ranks():
    _run(2)


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs >= 4 GPUs")
def test_allreduce_kernels_all_ranks():
    _run(min(8, torch.cuda.device_count()))
