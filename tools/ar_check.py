"""Multi-GPU check of the one-shot NVLink all-reduce (csrc/allreduce.cu) against ncclAllReduce, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29540 tools/ar_check.py

Checks: equal to the fp32 sum of the per-rank inputs rounded once (and within bf16 rounding of NCCL's result),
bit-identical on every rank, in-place, many back-to-back calls, CUDA-graph replay; then times both paths
(CUDA events over a graph of 50 calls) at the row-parallel message sizes of the BASELINE configs.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from iaas_sglang_b200 import tp  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    tp.init_distributed("nccl", dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    ar = tp.custom_all_reduce()
    assert ar is not None, "custom all-reduce did not initialise"
    ok = True
    for n in (8, 4096, 64 * 4096, 128 * 4096, 256 * 8192, 4 << 20):
        for dt in (torch.bfloat16, torch.float16):
            g = torch.Generator(device=dev).manual_seed(1000 * rank + n % 997)
            x = torch.randn(n, device=dev, dtype=dt, generator=g)
            parts = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(parts, x)
            acc = torch.zeros(n, device=dev, dtype=torch.float32)   # fp32 sum in rank order, rounded once
            for p in parts:
                acc += p.float()
            want = acc.to(dt)
            y = x.clone()
            ar.all_reduce_(y)
            same = torch.equal(y, want)
            z = x.clone()
            dist.all_reduce(z)
            close = (y.float() - z.float()).abs().max().item() <= 0.02 * max(1.0, z.float().abs().max().item())
            allr = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(allr, y)
            ident = all(torch.equal(allr[0], a) for a in allr)
            if rank == 0:
                print(f"n={n:8d} {str(dt):15s} exact_rank_order_sum={same} close_to_nccl={close} identical_on_all_ranks={ident}", flush=True)
            ok = ok and same and close and ident
    # many calls back to back + graph replay (the call counters live on the device)
    x = torch.full((128 * 4096,), float(rank + 1), device=dev, dtype=torch.bfloat16)
    want = float(sum(range(1, world + 1)))
    for _ in range(200):
        y = x.clone()
        ar.all_reduce_(y)
    torch.cuda.synchronize()
    ok = ok and bool((y == want).all().item())
    # one-shot and two-shot calls interleaved (they share the flags and the two data slots)
    small = torch.full((4096,), float(rank + 1), device=dev, dtype=torch.bfloat16)
    big = torch.full((2 << 20,), float(rank + 1), device=dev, dtype=torch.bfloat16)
    for i in range(60):
        a, c = small.clone(), big.clone()
        ar.all_reduce_(a)
        ar.all_reduce_(c)
        if i % 3 == 0:
            ar.all_reduce_(a.copy_(small))
    torch.cuda.synchronize()
    ok = ok and bool((a == want).all().item()) and bool((c == want).all().item())
    st = torch.cuda.Stream()
    buf = x.clone()
    with torch.cuda.stream(st):
        ar.all_reduce_(buf.copy_(x))
        torch.cuda.synchronize()
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph, stream=st):
            for _ in range(10):
                buf.copy_(x)
                ar.all_reduce_(buf)
    for _ in range(20):
        gph.replay()
    torch.cuda.synchronize()
    ok = ok and bool((buf == want).all().item()) and ar.error() == 0
    if rank == 0:
        print(f"back-to-back + graph replay: {'ok' if ok else 'FAILED'} (error word {ar.error()})", flush=True)

    # timing: graph of 50 calls, CUDA events, max over ranks
    def timeit(fn, t):
        s2 = torch.cuda.Stream()
        with torch.cuda.stream(s2):
            fn(t)
            torch.cuda.synchronize()
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2, stream=s2):
                for _ in range(50):
                    fn(t)
        g2.replay()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g2.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / 50], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() * 1e3

    cases = [("cfg2 tp%d 64xN x4096" % world, 64 * world * 4096), ("cfg3 128x5120", 128 * 5120), ("cfg4 256x8192", 256 * 8192)]
    if os.environ.get("AR_SWEEP"):
        cases = [(f"{kb} KB", kb * 512) for kb in (64, 256, 512, 1024, 2048, 4096, 8192)]
    for name, n in cases:
        t = torch.randn(n, device=dev, dtype=torch.bfloat16)
        us_c = timeit(lambda a: ar.all_reduce_(a), t)
        us_n = timeit(lambda a: dist.all_reduce(a), t)
        if rank == 0:
            print(f"{name:24s} {n * 2 / 1e6:6.2f} MB  one-shot {us_c:7.1f} us   nccl {us_n:7.1f} us", flush=True)
    dist.barrier()
    if rank == 0:
        print("AR_CHECK " + ("PASS" if ok else "FAIL"), flush=True)
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
