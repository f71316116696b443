"""Multi-GPU check of the NVLink / NVSwitch all-reduce kernels (csrc/allreduce.cu) against ncclAllReduce, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29540 tools/ar_check.py

Checks, for every kernel the context can run (one-shot, two-shot, row-layout one-shot, NVLS multicast):
  * equal to the fp32 sum of the per-rank inputs in rank order rounded once (peer-load kernels: bit-exact; NVLS: the
    switch's accumulation order is its own, so within one rounding of the half type) and close to NCCL's result;
  * bit-identical on every rank; in place; many calls back to back; CUDA-graph replay;
  * a MIXED sequence of sizes, kernels and grids on RANDOM data with one rank delayed (the data slots are shared by
    every call: a stale slot would show up here);
  * the fused all-reduce + residual-add + RMSNorm (+ FP8 quantisation) against all-reduce followed by
    ops.rmsnorm_quant_fp8: bit-identical with peer loads, in tolerance with NVLS.
Then times every kernel and NCCL (CUDA events over a graph of 50 calls, max over ranks) at the row-parallel message
sizes of the BASELINE configs (AR_SWEEP=1: 64 KB .. 8 MB).  Exit code 0 = all checks passed.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from iaas_sglang_b200 import ops, tp  # noqa: E402

MODE_NAMES = {1: "one-shot", 2: "two-shot", 3: "nvls", 4: "rows-pull", 5: "push-1", 6: "push-2"}


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    tp.init_distributed("nccl", dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    ar = tp.custom_all_reduce()
    assert ar is not None, "custom all-reduce did not initialise"
    modes = [1, 4] + ([2] if world > 2 else []) + ([3] if ar.multicast else []) + [5] + ([6] if world > 2 else [])
    if os.environ.get("AR_MODES"):
        modes = [int(m) for m in os.environ["AR_MODES"].split(",")]
    if rank == 0:
        print(f"world {world}: set-up {ar.setup}, multicast {ar.multicast}, kernels {[MODE_NAMES[m] for m in modes]}", flush=True)
    ok = True

    def say(msg):
        if rank == 0:
            print(msg, flush=True)

    def expected(x):
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x)
        acc = torch.zeros(x.shape, device=dev, dtype=torch.float32)   # fp32 sum in rank order, rounded once
        for p in parts:
            acc += p.float()
        return acc

    def identical_everywhere(y):
        allr = [torch.empty_like(y) for _ in range(world)]
        dist.all_gather(allr, y)
        return all(torch.equal(allr[0], a) for a in allr)

    # ---- 1. every kernel at a range of sizes and both dtypes ----
    for mode in modes:
        ar.set_mode(mode)
        for n in (8, 4096, 64 * 4096 + 8, 128 * 4096, 256 * 8192, 4 << 20):
            for dt in (torch.bfloat16, torch.float16):
                g = torch.Generator(device=dev).manual_seed(1000 * rank + n % 997)
                x = torch.randn(n, device=dev, dtype=dt, generator=g)
                acc = expected(x)
                want = acc.to(dt)
                y = x.clone()
                if mode == 5 and n * 2 * world > 2 * ar.max_bytes:
                    continue   # W copies of the message do not fit one receive buffer
                ar.all_reduce_(y)
                if mode == 3:   # one ulp of the half type around the exact sum
                    ulp = (want.float().abs() * (2.0 ** -7 if dt == torch.bfloat16 else 2.0 ** -10)).clamp_min(1e-6)
                    same = bool(((y.float() - acc).abs() <= ulp).all().item())
                else:
                    same = torch.equal(y, want)
                z = x.clone()
                dist.all_reduce(z)
                close = (y.float() - z.float()).abs().max().item() <= 0.02 * max(1.0, z.float().abs().max().item())
                ident = identical_everywhere(y)
                if not (same and close and ident):
                    say(f"FAIL {MODE_NAMES[mode]} n={n} {dt}: rank_order_sum={same} close_to_nccl={close} identical={ident}")
                ok = ok and same and close and ident
        say(f"{MODE_NAMES[mode]:10s}: sizes x dtypes {'ok' if ok else 'FAILED'}")

    # ---- 2. mixed sizes / kernels / grids on random data, one rank delayed ----
    sizes = [8 * 1024 // 2, 64 * 1024 // 2, 2 * 1024 * 1024 // 2, 24, 1024 * 1024 // 2, 4 * 1024 * 1024 // 2]
    gen = torch.Generator(device=dev).manual_seed(77 + rank)
    seq = []
    for i in range(36):
        n = sizes[i % len(sizes)]
        m = modes[i % len(modes)]
        if m == 5 and n * 2 * world > 2 * ar.max_bytes:
            m = 1
        seq.append((m, torch.randn(n, device=dev, dtype=torch.bfloat16, generator=gen)))
    wants = [expected(x) for _, x in seq]
    outs = []
    for i, (mode, x) in enumerate(seq):
        ar.set_mode(mode)
        if i % 5 == rank % 5:
            torch.cuda._sleep(3_000_000)   # ~1.5 ms: this rank arrives late, its peers are two calls ahead
        y = x.clone()
        ar.all_reduce_(y)
        outs.append(y)
    torch.cuda.synchronize()
    mixed = True
    for (mode, x), acc, y in zip(seq, wants, outs):
        tol = (acc.abs() * 2.0 ** -7).clamp_min(1e-6)
        good = bool(((y.float() - acc).abs() <= tol).all().item()) if mode == 3 else torch.equal(y, acc.to(torch.bfloat16))
        mixed = mixed and good and identical_everywhere(y)
    say(f"mixed sizes/kernels with a delayed rank: {'ok' if mixed else 'FAILED'}")
    ok = ok and mixed

    # ---- 3. back to back + CUDA-graph replay (the call counter lives on the device) ----
    ar.set_mode(-1)
    x = torch.full((128 * 4096,), float(rank + 1), device=dev, dtype=torch.bfloat16)
    want = float(sum(range(1, world + 1)))
    for _ in range(200):
        y = x.clone()
        ar.all_reduce_(y)
    torch.cuda.synchronize()
    ok = ok and bool((y == want).all().item())
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
    good = bool((buf == want).all().item()) and ar.error() == 0
    say(f"back-to-back + graph replay: {'ok' if good else 'FAILED'} (error word {ar.error()})")
    ok = ok and good

    # ---- 4. fused all-reduce + add + RMSNorm (+ FP8 quant) vs the two-kernel sequence ----
    for mode in [m for m in modes if m in (3, 4, 5, 6)]:
        ar.set_mode(mode)
        for (M, H) in ((64, 4096), (128, 5120), (256, 8192), (512, 4096), (37, 7168)):
            if mode == 5 and M * H * 2 * world > 2 * ar.max_bytes:
                continue
            g = torch.Generator(device=dev).manual_seed(5 + rank)
            x = (torch.randn(M, H, device=dev, generator=g) * 0.5).to(torch.bfloat16)
            gs = torch.Generator(device=dev).manual_seed(99)   # replicated tensors
            res0 = torch.randn(M, H, device=dev, generator=gs).to(torch.bfloat16)
            w = (torch.rand(H, device=dev, generator=gs) + 0.5).to(torch.bfloat16)
            # two kernels (the all-reduce of the same mode, then the fused add + norm + quant)
            h = x.clone(); r1 = res0.clone()
            ar.all_reduce_(h)
            q1, s1, y1 = ops.rmsnorm_quant_fp8(h, w, 1e-5, residual=r1, want_half=True)
            r2 = res0.clone()
            q2, s2, y2 = ar.all_reduce_add_rmsnorm(x.clone(), r2, w, 1e-5, quant_fp8=True, want_half=True)
            r3 = res0.clone()
            y3 = ar.all_reduce_add_rmsnorm(x.clone(), r3, w, 1e-5)
            torch.cuda.synchronize()
            if mode == 3:   # the switch's summation order may depend on the requesting rank: tolerance, not bits
                def rl2(a, b):
                    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-30)).item()
                same = (rl2(r2, r1) < 4e-3 and rl2(y2, y1) < 1e-2 and rl2(s2, s1) < 1e-2 and torch.equal(y2, y3) and torch.equal(r2, r3)
                        and rl2(q2.float(), q1.float()) < 5e-2)
            else:
                same = (torch.equal(r1, r2) and torch.equal(y1, y2) and torch.equal(q1.view(torch.uint8), q2.view(torch.uint8))
                        and torch.equal(s1, s2) and torch.equal(y1, y3) and torch.equal(r1, r3))
            ident = identical_everywhere(y2) and identical_everywhere(r2)
            if not (same and ident):
                say(f"FAIL fused {MODE_NAMES[mode]} M={M} H={H}: equal_to_two_kernels={same} identical={ident}")
            ok = ok and same and ident
        say(f"fused all-reduce + add + RMSNorm (+quant), {MODE_NAMES[mode]}: {'ok' if ok else 'FAILED'}")
    ar.set_mode(-1)

    # ---- timing: graph of 50 calls, CUDA events, max over ranks ----
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

    cases = [("cfg2 tp%d %dx4096" % (world, 64 * world), 64 * world * 4096), ("cfg3 128x5120", 128 * 5120),
             ("cfg5 256x7168", 256 * 7168), ("cfg4 256x8192", 256 * 8192)]
    if os.environ.get("AR_SWEEP"):
        cases = [(f"{kb} KB", kb * 512) for kb in (64, 256, 512, 1024, 2048, 4096, 8192)]
    say("message".ljust(26) + "".join(f"{MODE_NAMES[m]:>11s}" for m in modes) + f"{'nccl':>11s}{'policy':>11s}   (us)")
    for name, n in cases:
        t = torch.randn(n, device=dev, dtype=torch.bfloat16)
        cols = []
        for m in modes:
            ar.set_mode(m)
            cols.append(float("nan") if (m == 5 and n * 2 * world > 2 * ar.max_bytes) else timeit(lambda a: ar.all_reduce_(a), t))
        ar.set_mode(-1)
        us_n = timeit(lambda a: dist.all_reduce(a), t)
        us_p = timeit(lambda a: tp.tensor_model_parallel_all_reduce(a), t)
        say(f"{name:22s}{n * 2 / 1e6:5.2f}MB" + "".join(f"{c:11.1f}" for c in cols) + f"{us_n:11.1f}{us_p:11.1f}")
    # fused vs two kernels at the decode shapes
    for (M, H) in ((64 * world, 4096), (256, 8192), (128, 5120)):
        x = torch.randn(M, H, device=dev, dtype=torch.bfloat16)
        res = torch.randn(M, H, device=dev, dtype=torch.bfloat16)
        w = torch.ones(H, device=dev, dtype=torch.bfloat16)

        def two(a):
            tp.tensor_model_parallel_all_reduce(a)
            ops.rmsnorm_quant_fp8(a, w, 1e-5, residual=res)

        def fused(a):
            tp.tensor_model_parallel_fused_allreduce_rmsnorm(a, res, w, 1e-5, quant_fp8=True, want_half=False)
        line = f"all-reduce + add+norm+quant {M}x{H}: two kernels {timeit(two, x):7.1f} us   fused(policy) {timeit(fused, x):7.1f} us"
        for m in [m for m in modes if m in (3, 4, 5, 6)]:
            if m == 5 and M * H * 2 * world > 2 * ar.max_bytes:
                continue
            ar.set_mode(m)
            line += f"   {MODE_NAMES[m]} {timeit(lambda a: ar.all_reduce_add_rmsnorm(a, res, w, 1e-5, quant_fp8=True, want_half=False), x):6.1f}"
        ar.set_mode(-1)
        say(line)
    dist.barrier()
    say("AR_CHECK " + ("PASS" if ok else "FAIL"))
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
