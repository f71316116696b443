#!/usr/bin/env python
"""bench.py -- decode-step throughput of the hot path (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2|cfg3|cfg4|cfg5|cfg2-int8]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One "step" = one decode step of the whole layer stack for one batch: per layer RMSNorm(+residual) -> qkv linear ->
RoPE -> paged decode attention (fused KV append, reading req_to_token in place) -> o_proj -> [all-reduce + residual-add +
RMSNorm + FP8 quant in ONE kernel under TP] -> gate_up linear -> SiLU*mul -> down linear (-> fused all-reduce + norm of
the next layer).  Embedding lookup, lm_head and sampling are outside the hot path and are NOT in the step;
this is stated in the JSON (`config.excluded`).

Headline workload (`value`, every N) = BASELINE.json configs[1]: Llama-3.1-8B FP8 block-128 W8A8 decode, batch 64 x 4096
ctx per GPU, page_size 1, shuffled page table, 32 layers each with its own KV pool and weights, random init; N > 1 =
tensor parallel over the N ranks with batch 64*N ("weak": per-GPU attention bytes fixed).

The JSON line carries: value (device-timed, CUDA-graph replay, inputs resident), e2e (same step through the plugin API
with host buffers: pinned H2D of the step's inputs + D2H of the result inside the timed region), roofline (decode
attention kernel: CUDA events, algorithmic bytes / measured HBM peak), roofline_gemm (FP8 block-scaled GEMM TFLOP/s vs the
dense FP8 peak measured in the same run), tp_check (N > 1: the sharded stack against the unsharded one on the same
weights, identical on all ranks), extra.configs (the other north_star configs at this N: cfg 4 Llama-70B FP8 with shared
prefix, cfg 3 Qwen-32B AWQ, cfg 5 DeepSeek MLA + FP8 -- layer slices sized to fit), cpu_baseline + extra.cpu_ops (the
reference's own CPU kernels, oracle/_ref, on the box's host cores), clocks, gpu_launches.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

# ------------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------------
WORKLOADS = {
    # name: model shape + batch; quant in {"fp8", "awq", "int8"}; arch "gqa" (Llama/Qwen layer) or "mla" (DeepSeek)
    "cfg2": dict(model="Llama-3.1-8B", arch="gqa", layers=32, hidden=4096, hq=32, hkv=8, d=128, inter=14336, quant="fp8",
                 batch=64, ctx=4096, shared_prefix=0, theta=500000.0, scale_batch=True),
    "cfg2-int8": dict(model="Llama-3.1-8B", arch="gqa", layers=32, hidden=4096, hq=32, hkv=8, d=128, inter=14336,
                      quant="int8", batch=64, ctx=4096, shared_prefix=0, theta=500000.0, scale_batch=True),
    "cfg3": dict(model="Qwen2.5-32B", arch="gqa", layers=64, hidden=5120, hq=40, hkv=8, d=128, inter=27648, quant="awq",
                 batch=128, ctx=4096, shared_prefix=0, theta=1000000.0, scale_batch=False),
    "cfg4": dict(model="Llama-3.1-70B", arch="gqa", layers=80, hidden=8192, hq=64, hkv=8, d=128, inter=28672, quant="fp8",
                 batch=256, ctx=8192, shared_prefix=2048, theta=500000.0, scale_batch=False),
    # DeepSeek-V3 dense-MLP layer slice (SURVEY 8 cfg 5): MLA absorbed decode over the 576-wide latent pool + the FP8
    # block linears of SURVEY a11 (q_a+kv_a 2112x7168 replicated, q_b, o_proj, dense gate_up / down sharded)
    "cfg5": dict(model="DeepSeek-V3 layer slice", arch="mla", layers=61, hidden=7168, hq=128, hkv=1, d=576, dv=512,
                 q_lora=1536, kv_lora=512, rope=64, nope=128, vdim=128, inter=18432, quant="fp8", batch=256, ctx=16384,
                 shared_prefix=0, theta=10000.0, scale_batch=False),
}
DEFAULT_WORKLOAD = "cfg2"
GIB = 1 << 30


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (layer slice); 0 = the model's (or what fits)")
    ap.add_argument("--batch", type=int, default=0, help="override the per-job batch; 0 = the workload's")
    ap.add_argument("--tp", type=int, default=0, help="emulate TP-local shapes on ONE GPU (no collectives); 0 = --gpus")
    ap.add_argument("--extras", default="auto", help="auto | none | comma list of workloads measured as extra.configs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-gemm-roofline", action="store_true")
    ap.add_argument("--no-tp-check", action="store_true")
    ap.add_argument("--no-fused", action="store_true", help="separate rmsnorm / silu_and_mul / quantiser / all-reduce kernels (the reference's op sequence)")
    ap.add_argument("--kernels", action="store_true", help="also print per-kernel roofline lines to stderr")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md "clocks DURING the timed region")
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}

# ------------------------------------------------------------------------------------------------
# our arm: the decode layer stack on the B200 kernels, through the plugin / op surface
# ------------------------------------------------------------------------------------------------
class _Params:
    pass


def _rand_fp8_linear(n, k, dev, gen):
    """weight [N,K] e4m3 + weight_scale_inv [N/128,K/128] (fp8.py:236-300), random init."""
    lay = _Params()
    w = torch.randn(n, k, device=dev, dtype=torch.float32, generator=gen) * 100.0
    lay.weight = w.clamp_(-448, 448).to(torch.float8_e4m3fn)
    lay.weight_scale_inv = torch.rand((n + 127) // 128, (k + 127) // 128, device=dev, generator=gen) * 2e-4 + 1e-4
    return lay


def _rand_int8_linear(n, k, dev, gen):
    """weight stored as the [K,N] column-major view of int8 [N,K]; weight_scale [N,1] (w8a8_int8.py:88-123)."""
    lay = _Params()
    w = (torch.randn(n, k, device=dev, generator=gen) * 40).clamp_(-127, 127).round_().to(torch.int8)
    lay.weight = w.t()
    lay.weight_scale = torch.rand(n, 1, device=dev, generator=gen) * 4e-4 + 1e-4
    return lay


def _rand_awq_linear(n, k, dev, gen, dtype):
    """qweight [K,N/8] int32, qzeros [K/128,N/8] int32, scales [K/128,N] half (awq.py:104-184)."""
    lay = _Params()
    lay.qweight = torch.randint(0, 2 ** 31 - 1, (k, n // 8), device=dev, dtype=torch.int32, generator=gen)
    lay.qzeros = torch.randint(0, 2 ** 31 - 1, (k // 128, n // 8), device=dev, dtype=torch.int32, generator=gen)
    lay.scales = (torch.rand(k // 128, n, device=dev, generator=gen) * 4e-3 + 1e-3).to(dtype)
    return lay


def _slice_fp8(full, row_ranges, col_range):
    """TP shard of an FP8 block linear: rows (N) for column-parallel, columns (K) for row-parallel; every boundary is a
    multiple of 128, so the [N/128, K/128] scale grid shards with the weight."""
    lay = _Params()
    c0, c1 = col_range
    lay.weight = torch.cat([full.weight[r0:r1, c0:c1] for r0, r1 in row_ranges], dim=0).contiguous()
    lay.weight_scale_inv = torch.cat([full.weight_scale_inv[r0 // 128:(r1 + 127) // 128, c0 // 128:(c1 + 127) // 128]
                                      for r0, r1 in row_ranges], dim=0).contiguous()
    return lay


class StackBase:
    """What the measurement loop needs from a layer stack: static CUDA-graph buffers + pinned host copies of one step's
    inputs, eager forward(), capture / replay, and the algorithmic byte counts of SURVEY 8d."""

    def _init_step_buffers(self, batch, hidden, ctx, table, dev):
        from iaas_sglang_b200 import compat

        self.hidden_in = torch.zeros(batch, hidden, dtype=self.dtype, device=dev)
        self.positions = torch.zeros(batch, dtype=torch.int64, device=dev)
        self.rope_cs = None   # per-step rotary cos / sin rows (forward() fills it; None = RoPE as its own launch, same bits)
        self.seq_lens = torch.ones(batch, dtype=torch.int64, device=dev)
        self.req_pool_indices = torch.zeros(batch, dtype=torch.int64, device=dev)
        self.out_cache_loc = torch.zeros(batch, dtype=torch.int64, device=dev)
        self.fb = compat.ForwardBatch(compat.ForwardMode.DECODE, batch, req_pool_indices=self.req_pool_indices,
                                      seq_lens=self.seq_lens, out_cache_loc=self.out_cache_loc, seq_lens_sum=batch * ctx,
                                      seq_lens_cpu=self.seq_lens_cpu, positions=self.positions,
                                      req_to_token_pool=self.r2t_pool, token_to_kv_pool=self.kv_pool, attn_backend=self.backend)
        g = torch.Generator().manual_seed(4321)   # the same step inputs on every rank (replicated scheduler state)
        self.h_hidden = (torch.randn(batch, hidden, generator=g) * 0.5).to(self.dtype).pin_memory()
        self.h_positions = torch.full((batch,), ctx - 1, dtype=torch.int64).pin_memory()
        self.h_seq_lens = self.seq_lens_cpu.clone().pin_memory()
        self.h_rpi = torch.arange(batch, dtype=torch.int64).pin_memory()
        self.h_loc = table[:, ctx - 1].to(torch.int64).cpu().pin_memory()
        self.h_out = torch.empty(batch, hidden, dtype=self.dtype).pin_memory()
        self.out = None
        self.graph = None

    def h2d_bytes(self):
        return sum(t.numel() * t.element_size() for t in (self.h_hidden, self.h_positions, self.h_seq_lens, self.h_rpi, self.h_loc))

    def d2h_bytes(self):
        return self.h_out.numel() * self.h_out.element_size()

    def upload(self):
        self.hidden_in.copy_(self.h_hidden, non_blocking=True)
        self.positions.copy_(self.h_positions, non_blocking=True)
        self.seq_lens.copy_(self.h_seq_lens, non_blocking=True)
        self.req_pool_indices.copy_(self.h_rpi, non_blocking=True)
        self.out_cache_loc.copy_(self.h_loc, non_blocking=True)

    def download(self):
        self.h_out.copy_(self.out, non_blocking=True)

    def capture(self):
        from iaas_sglang_b200 import compat

        self.backend.init_cuda_graph_state(self.B, self.B)
        self.backend.init_forward_metadata_capture_cuda_graph(self.B, self.B, self.req_pool_indices, self.seq_lens, None,
                                                              compat.ForwardMode.DECODE, None)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self.out = self.forward()     # warm-up on the side stream (sizes workspaces, sets func attributes)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=s):
            self.out = self.forward()
        torch.cuda.synchronize()

    def replay(self):
        from iaas_sglang_b200 import compat

        self.backend.init_forward_metadata_replay_cuda_graph(self.B, self.req_pool_indices, self.seq_lens, self.B * self.wl["ctx"],
                                                             None, compat.ForwardMode.DECODE, None, seq_lens_cpu=self.seq_lens_cpu)
        self.graph.replay()

    def _page_table(self, batch, ctx, shared, dev):
        """shuffled slots (page_size 1, worst-case scatter); optional RadixAttention shared prefix: the first `shared`
        positions of every request point at the same slots (SURVEY 8d cfg 4)."""
        from iaas_sglang_b200 import compat

        gen_shared = torch.Generator(device=dev)   # scheduler state is replicated on every TP rank
        gen_shared.manual_seed(1234)
        unique = batch * (ctx - shared) + shared
        perm = torch.randperm(unique, device=dev, generator=gen_shared) + 1  # slot 0 reserved (memory_pool.py:236-239)
        self.r2t_pool = compat.ReqToTokenPool(batch, ctx, dev)
        table = self.r2t_pool.req_to_token
        if shared:
            table[:, :shared] = perm[:shared].to(torch.int32)[None, :]
        table[:, shared:] = perm[shared:].view(batch, ctx - shared).to(torch.int32)
        self.seq_lens_cpu = torch.full((batch,), ctx, dtype=torch.int64)
        self.T = unique
        return table

    # fused (all-reduce +) residual-add + RMSNorm (+ FP8 group quant): ONE kernel; x is an un-reduced row-parallel
    # partial when reduce=True (layers/communicator.py + flashinfer_comm_fusion.py:122 in the reference)
    def _norm(self, x, w, residual, reduce, quant):
        from iaas_sglang_b200 import ops, tp

        if reduce and self.collectives:
            if self.no_fused:
                x = tp.tensor_model_parallel_all_reduce(x)
            else:
                return tp.tensor_model_parallel_fused_allreduce_rmsnorm(x, residual, w, 1e-5, quant_fp8=quant, want_half=False)
        if quant:
            return ops.rmsnorm_quant_fp8(x, w, 1e-5, residual=residual)
        return ops.rmsnorm(x, w, 1e-5, residual=residual)


class DecodeStack(StackBase):
    """L decoder layers (python/sglang/srt/models/llama.py LlamaDecoderLayer shape) sharded for TP rank `rank`.
    shared_weights=True: every rank draws the FULL model from one seed and keeps its shard (tp_check compares the
    sharded stack with the unsharded one); otherwise each rank draws only its own shard (same shapes, less memory)."""

    def __init__(self, wl: dict, layers: int, batch: int, tp: int, rank: int, dev: torch.device, collectives: bool,
                 shared_weights: bool = False, ctx: int = 0):
        from iaas_sglang_b200 import compat, linear
        from iaas_sglang_b200.backend import B200AttnBackend

        self.wl, self.L, self.B, self.tp, self.dev = dict(wl), layers, batch, tp, dev
        if ctx:
            self.wl["ctx"] = ctx
        wl = self.wl
        self.collectives = collectives
        self.no_fused = False
        self.probe = None
        self.dtype = torch.bfloat16
        hq, hkv, d, hidden, inter = wl["hq"], wl["hkv"], wl["d"], wl["hidden"], wl["inter"]
        assert hq % tp == 0 and inter % tp == 0 and (hkv % tp == 0 or tp % hkv == 0)
        self.hq, self.hkv, self.d, self.hidden = hq // tp, max(1, hkv // tp), d, hidden
        self.inter = inter // tp
        ctx, shared = wl["ctx"], wl["shared_prefix"]
        gen = torch.Generator(device=dev)
        gen.manual_seed(1234 if shared_weights else 1234 + rank)

        table = self._page_table(batch, ctx, shared, dev)
        self.kv_pool = compat.MHATokenToKVPool(self.T, 1, self.dtype, self.hkv, d, layers, dev)
        kv_rank = (rank * hkv) // tp if hkv >= tp else rank // (tp // hkv)   # first kv head of this rank
        for l in range(layers):
            if shared_weights:   # the full pool from the shared seed, this rank's heads kept
                for buf in (self.kv_pool.k_buffer[l], self.kv_pool.v_buffer[l]):
                    full = torch.empty(buf.shape[0], hkv, d, dtype=self.dtype, device=dev).normal_(generator=gen)
                    buf.copy_(full[:, kv_rank:kv_rank + self.hkv])
            else:
                self.kv_pool.k_buffer[l].normal_(generator=gen)
                self.kv_pool.v_buffer[l].normal_(generator=gen)

        runner = compat.ModelRunnerStub(compat.ModelConfigStub(hq, hkv, d, ctx), self.r2t_pool, self.kv_pool, device=dev, tp_size=tp)
        # Cascade decode (shared prefix attended once per batch) is measured BESIDE the plain path for workloads with a
        # RadixAttention shared prefix (run_extra: roofline.cascade): on B200 the plain kernel already serves a 2048-token
        # shared prefix from the 126 MB L2, so cascade saves L2 traffic, not HBM bytes, and is not the faster plan there.
        self.cascade = bool(shared) and os.environ.get("B200_BENCH_CASCADE", "0") == "1"
        self.runner = runner
        self.backend = B200AttnBackend(runner, cascade=self.cascade)
        self.attn = [compat.RadixAttention(self.hq, d, d ** -0.5, self.hkv, l) for l in range(layers)]

        # ---- weights ----
        q = wl["quant"]
        qd, kvd, it = self.hq * d, self.hkv * d, self.inter
        self.qkv_dim = qd + 2 * kvd
        if q == "fp8":
            self.method = linear.Fp8BlockLinearMethod()
        elif q == "int8":
            self.method = linear.W8A8Int8LinearMethod()
        elif q == "awq":
            self.method = linear.AWQLinearMethod()
        else:
            raise SystemExit(f"bench: quant {q} not benchable")

        def mk(n, k):
            if q == "fp8":
                return _rand_fp8_linear(n, k, dev, gen)
            if q == "int8":
                return _rand_int8_linear(n, k, dev, gen)
            return _rand_awq_linear(n, k, dev, gen, self.dtype)

        self.w_qkv, self.w_o, self.w_gate_up, self.w_down = [], [], [], []
        for _ in range(layers):
            if shared_weights:
                assert q == "fp8" and hkv % tp == 0
                Q, KV = hq * d, hkv * d
                f = _rand_fp8_linear(Q + 2 * KV, hidden, dev, gen)
                self.w_qkv.append(_slice_fp8(f, [(rank * qd, (rank + 1) * qd), (Q + rank * kvd, Q + (rank + 1) * kvd),
                                                 (Q + KV + rank * kvd, Q + KV + (rank + 1) * kvd)], (0, hidden)))
                f = _rand_fp8_linear(hidden, Q, dev, gen)
                self.w_o.append(_slice_fp8(f, [(0, hidden)], (rank * qd, (rank + 1) * qd)))
                f = _rand_fp8_linear(2 * inter, hidden, dev, gen)
                self.w_gate_up.append(_slice_fp8(f, [(rank * it, (rank + 1) * it), (inter + rank * it, inter + (rank + 1) * it)], (0, hidden)))
                f = _rand_fp8_linear(hidden, inter, dev, gen)
                self.w_down.append(_slice_fp8(f, [(0, hidden)], (rank * it, (rank + 1) * it)))
                del f
            else:
                self.w_qkv.append(mk(self.qkv_dim, hidden))
                self.w_o.append(mk(hidden, qd))
                self.w_gate_up.append(mk(2 * it, hidden))
                self.w_down.append(mk(hidden, it))
        self.norm1 = [torch.ones(hidden, dtype=self.dtype, device=dev) for _ in range(layers)]
        self.norm2 = [torch.ones(hidden, dtype=self.dtype, device=dev) for _ in range(layers)]
        self.norm_f = torch.ones(hidden, dtype=self.dtype, device=dev)
        self._init_step_buffers(batch, hidden, ctx, table, dev)

    def attn_block(self, l, x, fused):
        """qkv linear -> RoPE -> paged decode attention (fused KV append) -> o_proj: this rank's row-parallel partial"""
        from iaas_sglang_b200 import ops

        d = self.d
        if fused:   # RoPE in the qkv GEMM's epilogue where the decode-shape kernel serves the shape (same bits, one launch less)
            qkv = self.method.apply_quantized(self.w_qkv[l], x[0], x[1], self.dtype,
                                              rope=(self.positions, (self.hq + self.hkv) * d, d, self.wl["theta"], self.rope_cs))
            q, k, v = qkv.split([self.hq * d, self.hkv * d, self.hkv * d], dim=-1)   # strided views, no copies
        else:
            qkv = self.method.apply(self.w_qkv[l], x)
            q, k, v = qkv.split([self.hq * d, self.hkv * d, self.hkv * d], dim=-1)
            ops.rope_neox_(q, k, self.positions, d, self.wl["theta"])
        if fused:   # the attention merge kernel emits the group-128 quantisation of its output itself
            a = self.attn[l](q, k, v, self.fb, quant_fp8=True)
            return self.method.apply_quantized(self.w_o[l], a[0], a[1], self.dtype)
        return self.method.apply(self.w_o[l], self.attn[l](q, k, v, self.fb))

    def mlp_block(self, l, x, fused):
        """gate_up linear -> SiLU*mul -> down linear: this rank's row-parallel partial"""
        from iaas_sglang_b200 import ops

        if fused:
            gu = self.method.apply_quantized(self.w_gate_up[l], x[0], x[1], self.dtype)
            act = ops.silu_and_mul_quant_fp8(gu)
            return self.method.apply_quantized(self.w_down[l], act[0], act[1], self.dtype)
        return self.method.apply(self.w_down[l], ops.silu_and_mul(self.method.apply(self.w_gate_up[l], x)))

    def forward(self) -> torch.Tensor:
        """One decode step over the static buffers (eager; also the body that gets graph-captured)."""
        from iaas_sglang_b200 import ops

        h = self.hidden_in
        residual = torch.zeros_like(h)
        fused = hasattr(self.method, "apply_quantized") and not self.no_fused
        # the rotary cos / sin rows of this step's positions, once for all layers (the qkv GEMM's epilogue rotates with them)
        self.rope_cs = ops.rope_cos_sin(self.positions, self.d, self.wl["theta"]) if fused and self.d == 128 else None
        for l in range(self.L):
            # FP8: the norm / activation / all-reduce kernels emit the group-128 quantisation themselves (bit-identical)
            x = self._norm(h, self.norm1[l], residual, reduce=l > 0, quant=fused)   # residual += [allreduce](h) ; x = norm(residual)
            h = self.attn_block(l, x, fused)
            x = self._norm(h, self.norm2[l], residual, reduce=True, quant=fused)
            if l == 0 and self.probe is not None:
                self.probe["residual_after_first_reduce"] = residual.clone()
            h = self.mlp_block(l, x, fused)
        return self._norm(h, self.norm_f, residual, reduce=True, quant=False)

    # ---- algorithmic work ----
    def attn_bytes_per_layer(self, unique: bool = False):
        B, ctx, d = self.B, self.wl["ctx"], self.d
        tokens = self.T if unique else B * ctx
        return tokens * self.hkv * 2 * d * 2 + B * ctx * 4 + B * self.hq * 2 * d * 2 + B * self.hkv * 2 * d * 2 * 2

    def weight_bytes_per_layer(self):
        n = self.qkv_dim * self.hidden + self.hidden * self.hq * self.d + 2 * self.inter * self.hidden + self.hidden * self.inter
        return n * (0.5 if self.wl["quant"] == "awq" else 1.0)

    def attn_inputs(self):
        dev = self.dev
        return (torch.randn(self.B, self.hq * self.d, device=dev, dtype=self.dtype),
                torch.randn(self.B, self.hkv * self.d, device=dev, dtype=self.dtype),
                torch.randn(self.B, self.hkv * self.d, device=dev, dtype=self.dtype), {})

    attn_kernel = "decode_attn_mma_kernel (+ split merge)"


class MlaDecodeStack(StackBase):
    """DeepSeek-V3 dense-MLP decoder layers with MLA in the absorbed decode form (models/deepseek_v2.py
    forward_absorb, :1235-1304), TP-local: the latent pool [T, 1, 576] is replicated, the 128 heads shard.
    The FP8 block linears are SURVEY a11's cfg 5 shapes.  The two absorb bmm (q_nope x W_UK, o x W_UV; bf16, 8f rank 4,
    not a row of 8a) are torch.bmm.  q_b_proj's rows are laid out [all heads' nope | all heads' rope] (a row
    permutation of a random-init weight) so that the rope part is one column slice; rope is the neox kernel
    (DeepSeek's is the interleaved form: same bytes, same flops)."""

    attn_kernel = "decode_mla_mma_kernel (+ split merge)"

    def __init__(self, wl: dict, layers: int, batch: int, tp: int, rank: int, dev: torch.device, collectives: bool, ctx: int = 0):
        from iaas_sglang_b200 import compat, linear
        from iaas_sglang_b200.backend import B200AttnBackend

        self.wl, self.L, self.B, self.tp, self.dev = dict(wl), layers, batch, tp, dev
        if ctx:
            self.wl["ctx"] = ctx
        wl = self.wl
        self.collectives, self.no_fused, self.dtype = collectives, False, torch.bfloat16
        hidden, H = wl["hidden"], wl["hq"] // tp
        self.H, self.hidden, self.inter = H, hidden, wl["inter"] // tp
        self.hq, self.hkv, self.d = H, 1, 576
        ql, kvl, rope, nope, vd = wl["q_lora"], wl["kv_lora"], wl["rope"], wl["nope"], wl["vdim"]
        ctx = wl["ctx"]
        gen = torch.Generator(device=dev)
        gen.manual_seed(1234 + rank)
        table = self._page_table(batch, ctx, 0, dev)
        self.kv_pool = compat.MLATokenToKVPool(self.T, 1, self.dtype, kvl, rope, layers, dev)
        for l in range(layers):
            self.kv_pool.kv_buffer[l].normal_(generator=gen)
        mc = compat.ModelConfigStub(wl["hq"], 1, kvl + rope, ctx, v_head_dim=kvl, attention_arch_is_mla=True, kv_lora_rank=kvl,
                                    qk_rope_head_dim=rope)
        runner = compat.ModelRunnerStub(mc, self.r2t_pool, self.kv_pool, device=dev, tp_size=tp)
        self.backend = B200AttnBackend(runner)
        scale = (nope + rope) ** -0.5
        self.attn = [compat.RadixAttention(H, kvl + rope, scale, 1, l, v_head_dim=kvl) for l in range(layers)]
        self.method = linear.Fp8BlockLinearMethod()
        mk = lambda n, k: _rand_fp8_linear(n, k, dev, gen)  # noqa: E731
        self.a_dim = ql + kvl + rope                                  # 2112: q_a (1536) | kv_a latent (512) | k rope (64)
        self.w_a = [mk(self.a_dim, hidden) for _ in range(layers)]                    # replicated
        self.w_qb = [mk(H * (nope + rope), ql) for _ in range(layers)]                # column-parallel
        self.w_uk = [(torch.randn(H, nope, kvl, device=dev, generator=gen) * 0.05).to(self.dtype) for _ in range(layers)]
        self.w_uv = [(torch.randn(H, kvl, vd, device=dev, generator=gen) * 0.05).to(self.dtype) for _ in range(layers)]
        self.w_o = [mk(hidden, H * vd) for _ in range(layers)]                        # row-parallel
        self.w_gate_up = [mk(2 * self.inter, hidden) for _ in range(layers)]
        self.w_down = [mk(hidden, self.inter) for _ in range(layers)]
        ones = lambda n: torch.ones(n, dtype=self.dtype, device=dev)  # noqa: E731
        self.norm1 = [ones(hidden) for _ in range(layers)]
        self.norm2 = [ones(hidden) for _ in range(layers)]
        self.norm_qa = [ones(ql) for _ in range(layers)]
        self.norm_kv = [ones(kvl) for _ in range(layers)]
        self.norm_f = ones(hidden)
        self.dims = (ql, kvl, rope, nope, vd)
        self._init_step_buffers(batch, hidden, ctx, table, dev)

    def forward(self) -> torch.Tensor:
        from iaas_sglang_b200 import ops

        ql, kvl, rope, nope, vd = self.dims
        B, H, m, fb = self.B, self.H, self.method, self.fb
        h = self.hidden_in
        residual = torch.zeros_like(h)
        for l in range(self.L):
            x = self._norm(h, self.norm1[l], residual, reduce=l > 0, quant=True)
            a = m.apply_quantized(self.w_a[l], x[0], x[1], self.dtype)                       # [B, 2112]
            q_a, kv_a, k_pe = a[:, :ql], a[:, ql:ql + kvl], a[:, ql + kvl:]
            xq = ops.rmsnorm_quant_fp8(q_a.contiguous(), self.norm_qa[l], 1e-6)
            qb = m.apply_quantized(self.w_qb[l], xq[0], xq[1], self.dtype)                   # [B, H*128 | H*64]
            kv_n = ops.rmsnorm(kv_a.contiguous(), self.norm_kv[l], 1e-6)                     # [B, 512]
            q_pe = qb[:, H * nope:]
            ops.rope_neox_(q_pe, k_pe, self.positions, rope, self.wl["theta"])
            q_lat = torch.bmm(qb[:, :H * nope].view(B, H, nope).transpose(0, 1), self.w_uk[l]).transpose(0, 1)   # [B, H, 512]
            o_lat = self.attn[l](q_lat.reshape(B, H * kvl), kv_n.view(B, 1, kvl), kv_n.view(B, 1, kvl), fb,
                                 q_rope=q_pe.view(B, H, rope), k_rope=k_pe.reshape(B, 1, rope))
            o = torch.bmm(o_lat.view(B, H, kvl).transpose(0, 1), self.w_uv[l]).transpose(0, 1).reshape(B, H * vd)
            h = m.apply(self.w_o[l], o)                                                      # row-parallel partial
            x = self._norm(h, self.norm2[l], residual, reduce=True, quant=True)
            gu = m.apply_quantized(self.w_gate_up[l], x[0], x[1], self.dtype)
            act = ops.silu_and_mul_quant_fp8(gu)
            h = m.apply_quantized(self.w_down[l], act[0], act[1], self.dtype)
        return self._norm(h, self.norm_f, residual, reduce=True, quant=False)

    def attn_bytes_per_layer(self, unique: bool = False):
        B, ctx = self.B, self.wl["ctx"]
        return B * ctx * (1152 + 4) + B * self.H * (576 + 512) * 2 + B * 1152 * 2

    def weight_bytes_per_layer(self):
        ql, kvl, rope, nope, vd = self.dims
        H = self.H
        fp8 = self.a_dim * self.hidden + H * (nope + rope) * ql + self.hidden * H * vd + 3 * self.inter * self.hidden
        return fp8 + 2 * (H * nope * kvl + H * kvl * vd)

    def attn_inputs(self):
        dev, (ql, kvl, rope, nope, vd) = self.dev, self.dims
        return (torch.randn(self.B, self.H * kvl, device=dev, dtype=self.dtype),
                torch.randn(self.B, 1, kvl, device=dev, dtype=self.dtype), torch.randn(self.B, 1, kvl, device=dev, dtype=self.dtype),
                dict(q_rope=torch.randn(self.B, self.H, rope, device=dev, dtype=self.dtype),
                     k_rope=torch.randn(self.B, 1, rope, device=dev, dtype=self.dtype)))


def make_stack(wl, layers, batch, tp, rank, dev, collectives, **kw):
    cls = MlaDecodeStack if wl.get("arch") == "mla" else DecodeStack
    return cls(wl, layers, batch, tp, rank, dev, collectives, **kw)


def auto_layers(wl: dict, tp: int, batch: int, cap: int, budget_gb: float = 110.0) -> int:
    """Layer slice that fits one GPU: KV pool + weights per layer against `budget_gb`, at most `cap` layers."""
    if wl.get("arch") == "mla":
        per = batch * wl["ctx"] * 1152 + 1.3e8
    else:
        hkv = max(1, wl["hkv"] // tp)
        unique = batch * (wl["ctx"] - wl["shared_prefix"]) + wl["shared_prefix"]
        w = ((wl["hq"] + 2 * wl["hkv"]) * wl["d"] * wl["hidden"] + wl["hidden"] * wl["hq"] * wl["d"] + 3 * wl["inter"] * wl["hidden"]) / tp
        per = unique * hkv * wl["d"] * 2 * 2 + w * (0.6 if wl["quant"] == "awq" else 1.0)
    return int(max(1, min(cap, wl["layers"], budget_gb * 1e9 // per)))


# ------------------------------------------------------------------------------------------------
# CPU leg: the reference's own CPU kernels (oracle/_ref) or the plain-C port, on a bounded sample
# ------------------------------------------------------------------------------------------------
class CpuLayerSample:
    """ONE decoder layer's hot path (decode_attention_cpu + the four quantised linears) at the workload's full batch
    and context, on host cores.  tokens/s = batch / (layer time x n_layers): a bounded sample (1 of L layers per
    step), extrapolated linearly over the identical layers -- stated in `sample`."""

    def __init__(self, wl: dict, batch: int, threads: int):
        from oracle import ref_ops

        self.wl, self.B = wl, batch
        self.use_ref = ref_ops.available()
        self.kind = "reference" if self.use_ref else "port"
        torch.set_num_threads(threads)
        self.threads = threads if self.use_ref else min(threads, int(os.environ.get("OMP_NUM_THREADS", threads)))
        g = torch.Generator().manual_seed(1234)
        hq, hkv, d, hidden, inter, ctx = wl["hq"], wl["hkv"], wl["d"], wl["hidden"], wl["inter"], wl["ctx"]
        T = batch * ctx
        bf = torch.bfloat16
        # bf16 noise straight from random 16-bit patterns would hold NaNs; draw small-int noise and scale instead
        self.k_buf = (torch.randint(-64, 64, (T + 1, hkv, d), generator=g, dtype=torch.int8).to(bf)) / 32
        self.v_buf = (torch.randint(-64, 64, (T + 1, hkv, d), generator=g, dtype=torch.int8).to(bf)) / 32
        self.r2t = (torch.randperm(T, generator=g) + 1).view(batch, ctx).to(torch.int32)
        self.seq_lens = torch.full((batch,), ctx, dtype=torch.int64)
        self.rpi = torch.arange(batch, dtype=torch.int64)
        self.loc = self.r2t[:, ctx - 1].to(torch.int64)
        self.logits = torch.zeros(batch, hq, 8, d + 1, dtype=torch.float32)   # intel_amx_backend.py:36-46
        self.x = (torch.randn(batch, hidden, generator=g) * 0.5).to(bf)
        self.shapes = [((hq + 2 * hkv) * d, hidden), (hidden, hq * d), (2 * inter, hidden), (hidden, inter)]
        self.ws = []
        for n, k in self.shapes:
            w = torch.randint(-100, 100, (n, k), generator=g, dtype=torch.int8).to(torch.float32).to(torch.float8_e4m3fn)
            s = torch.rand(n // 128, k // 128, generator=g) * 2e-4 + 1e-4
            if self.use_ref:   # prepacked once, like the reference's weight loader (convert_weight_packed, is_vnni=True)
                w = ref_ops.ops().convert_weight_packed(w)
            self.ws.append((w, s))
        self.o = torch.zeros(batch, hq, d, dtype=bf)
        self.quant = wl["quant"]

    def _linear(self, x, i):
        w, s = self.ws[i]
        if self.use_ref:
            from oracle import ref_ops
            return ref_ops.ops().fp8_scaled_mm_cpu(x, w, s, [128, 128], None, torch.bfloat16, True)
        from oracle import cpu_oracle as orc
        return orc.fp8_w8a16_mm(x, w, s, (128, 128), None, torch.bfloat16)

    def layer(self):
        wl = self.wl
        hq, hkv, d = wl["hq"], wl["hkv"], wl["d"]
        qkv = self._linear(self.x, 0)
        q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
        q, k, v = q.reshape(-1, hq, d), k.reshape(-1, hkv, d), v.reshape(-1, hkv, d)
        if self.use_ref:
            from oracle import ref_ops
            ref_ops.decode_attention_cpu(q, self.k_buf, self.v_buf, self.o, k, v, self.loc, self.logits, self.r2t, self.rpi,
                                         self.seq_lens, d ** -0.5, 0.0)
        else:
            from oracle import cpu_oracle as orc
            orc.decode_attention(q, self.k_buf, self.v_buf, self.o, k, v, self.loc, self.logits, self.r2t, self.rpi,
                                 self.seq_lens, d ** -0.5, 0.0)
        h = self._linear(self.o.view(-1, hq * d), 1)
        gu = self._linear(h, 2)
        dd = gu.shape[-1] // 2
        act = torch.nn.functional.silu(gu[:, :dd]) * gu[:, dd:]
        return self._linear(act, 3)

    def describe(self, n_layers):
        return (f"1 of {n_layers} layers per step at the full batch {self.B} x {self.wl['ctx']} ctx "
                f"(decode_attention_cpu + 4 fp8 block linears via fp8_scaled_mm_cpu, W8A16 as the reference's CPU path is); "
                f"tokens/s = batch / (layer time x {n_layers})")


def time_cpu_layers(sample: CpuLayerSample, warmup: int, steps: int, n_layers: int, budget_s: float = 1e9):
    for _ in range(max(1, warmup)):
        sample.layer()
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        sample.layer()
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = (time.perf_counter() - t0) / done
    return sample.B / (dt * n_layers), dt * 1e3, done


def run_reference_arm(args, wl, n_layers):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref: decode_attention_cpu +
    fp8_scaled_mm_cpu) on this box's host cores.  A step = one bounded sample (ONE layer at the full batch), the line's
    ms_per_step / tokens/s are that sample extrapolated over the identical layers (stated in `sample`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    batch = args.batch or wl["batch"] * (max(1, args.gpus) if wl.get("scale_batch") else 1)
    if wl["quant"] != "fp8" or wl.get("arch") == "mla":
        emit({"impl": "reference", "unavailable": f"CPU leg implemented for the GQA fp8 workloads, not {args.workload}"})
        return
    sample = CpuLayerSample(wl, batch, cores)
    val, ms, done = time_cpu_layers(sample, args.warmup, args.steps, n_layers)
    line = {
        "impl": "reference", "metric": "decode tok/s", "value": round(val, 3), "unit": "tok/s", "n_gpus": args.gpus,
        "steps": done, "warmup": args.warmup, "ms_per_step": round(ms * n_layers, 3), "ms_per_sampled_layer": round(ms, 3),
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16 activations, fp8 e4m3 weights (W8A16 on CPU)", "data": "synthetic",
        "config": workload_config(wl, args.workload, n_layers, batch, 1),
        "cpu_baseline": {"value": round(val, 3), "unit": "tok/s", "cores": sample.threads, "kind": sample.kind,
                         "sample": sample.describe(n_layers)},
        "e2e": {"value": round(val, 3), "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(wl, name, n_layers, batch, tp):
    return {
        "workload": f"{wl['model']} {wl['quant']} decode, batch {batch} x {wl['ctx']} ctx, page_size 1 (shuffled page table)"
                    + (f", shared prefix {wl['shared_prefix']}" if wl["shared_prefix"] else ""),
        "name": name, "layers": n_layers, "model_layers": wl["layers"], "global_batch": batch, "ctx": wl["ctx"], "parallelism": f"tp{tp}",
        "l2": "inputs larger than L2 (KV >= 0.5 GB and weights >= 60 MB per layer per GPU, each layer its own buffers)",
        "excluded": "embedding lookup, lm_head, sampling (outside the hot path, SURVEY 8d)",
    }


# ------------------------------------------------------------------------------------------------
# measurement helpers
# ------------------------------------------------------------------------------------------------
def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                p = json.load(f)
            return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


_OUT_FD = 1


def claim_stdout():
    """stdout carries exactly ONE line, the JSON: whatever else writes to fd 1 (NCCL's version banner under torchrun)
    is sent to stderr for the rest of the run."""
    global _OUT_FD
    sys.stdout.flush()
    _OUT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    sys.stdout.flush()
    os.write(_OUT_FD, (json.dumps(line) + "\n").encode())


def log(*a):
    print(*a, file=sys.stderr, flush=True)


class Ctx:
    """rank / world plumbing shared by the measurement helpers"""

    def __init__(self, world, rank, dev):
        self.world, self.rank, self.dev = world, rank, dev

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        if self.world == 1:
            return list(vals)
        import torch.distributed as dist

        t = torch.tensor(list(vals), device=self.dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()


def measure_stack(cx: Ctx, stack, steps: int, warmup: int, e2e: bool, lib):
    """value: CUDA-graph replay, inputs resident, CUDA events, max over ranks.  e2e: pinned H2D -> replay -> D2H, stream
    sync per step, wall clock between barriers.  roofline: the attention kernel (+ merge) alone over every layer's pool."""
    stack.upload()
    n0 = lib.b200_launch_count()
    stack.capture()
    launches_per_step = (lib.b200_launch_count() - n0) // 2   # one eager warm-up + one captured pass
    for _ in range(max(3, warmup)):
        stack.replay()
    cx.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cx.barrier()
    ev0.record()
    for _ in range(steps):
        stack.replay()
    ev1.record()
    cx.barrier()
    ms = ev0.elapsed_time(ev1) / steps
    e2e_ms = 0.0
    if e2e:
        for _ in range(3):
            stack.upload(); stack.replay(); stack.download()
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            stack.upload()
            stack.replay()
            stack.download()
            torch.cuda.current_stream().synchronize()   # the caller needs the result before the next step
        cx.barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
    # ---- the dominant kernel alone (paged decode attention), CUDA events on the launching stream ----
    stack.backend.init_forward_metadata(stack.fb)
    q, k, v, kw = stack.attn_inputs()
    L = stack.L
    for l in range(min(3, L)):
        stack.attn[l](q, k, v, stack.fb, **kw)
    reps = max(1, 64 // L)
    cx.barrier()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record()
    for _ in range(reps):
        for l in range(L):
            stack.attn[l](q, k, v, stack.fb, **kw)
    a1.record()
    cx.barrier()
    attn_ms = a0.elapsed_time(a1) / (reps * L)
    ms, e2e_ms, attn_ms = cx.max_over_ranks([ms, e2e_ms, attn_ms])
    return dict(ms=ms, e2e_ms=e2e_ms if e2e else None, attn_ms=attn_ms, launches_per_step=launches_per_step)


def measure_cascade_attention(cx: Ctx, stack, plain_ms: float):
    """the same attention calls through a second backend with cascade decode on (shared prefix once per batch + suffix +
    LSE merge), timed like the plain ones"""
    from iaas_sglang_b200 import compat
    from iaas_sglang_b200.backend import B200AttnBackend

    be = B200AttnBackend(stack.runner, cascade=True)
    fb = compat.ForwardBatch(compat.ForwardMode.DECODE, stack.B, req_pool_indices=stack.req_pool_indices, seq_lens=stack.seq_lens,
                             out_cache_loc=stack.out_cache_loc, seq_lens_sum=stack.B * stack.wl["ctx"], seq_lens_cpu=stack.seq_lens_cpu,
                             positions=stack.positions, req_to_token_pool=stack.r2t_pool, token_to_kv_pool=stack.kv_pool, attn_backend=be)
    be.init_forward_metadata(fb)
    q, k, v, kw = stack.attn_inputs()
    L = stack.L
    for l in range(min(2, L)):
        stack.attn[l](q, k, v, fb, **kw)
    cx.barrier()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record()
    for l in range(L):
        stack.attn[l](q, k, v, fb, **kw)
    a1.record()
    cx.barrier()
    ms = cx.max_over_ranks([a0.elapsed_time(a1) / L])[0]
    return {"shared_prefix_len_detected": int(be.cascade_state.prefix_len.item()), "ms_per_launch": round(ms, 4),
            "vs_plain": round(plain_ms / ms, 3),
            "note": "opt-in (B200_CASCADE=1 / B200AttnBackend(cascade=True)); same result as the plain path (tests/test_gpu_cascade.py)"}


def roofline_of(stack, attn_ms, peak, peak_src, traffic=None, traffic_src=None):
    nbytes = stack.attn_bytes_per_layer()
    achieved = nbytes / (attn_ms * 1e-3) / 1e9
    r = {"kernel": stack.attn_kernel, "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
         "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src, "ms_per_launch": round(attn_ms, 4),
         "algorithmic_bytes_per_launch": int(nbytes), "frac_of_nominal_8TBs": round(achieved / 8000.0, 4)}
    if traffic_src:
        r["traffic_source"] = traffic_src
    uniq = stack.attn_bytes_per_layer(unique=True)
    if uniq != nbytes:   # shared prefix: logical bytes are the headline, unique bytes reported beside them
        r["unique_bytes_per_launch"] = int(uniq)
        r["frac_of_unique_bytes"] = round(uniq / (attn_ms * 1e-3) / 1e9 / peak, 4)
    return r


def step_roofline_of(stack, ms, peak):
    step_bytes = stack.L * (stack.attn_bytes_per_layer() + stack.weight_bytes_per_layer())
    return {"algorithmic_bytes_per_step_per_gpu": int(step_bytes), "hbm_time_ms_at_peak": round(step_bytes / peak / 1e6, 4),
            "frac": round(step_bytes / peak / 1e6 / ms, 4)}


def time_allreduce(cx: Ctx, rows: int, hidden: int):
    """the row-parallel reduction of this workload alone: plain all-reduce and the fused all-reduce + add + RMSNorm + FP8
    quant, CUDA-graph replay of 20 calls, max over ranks; algorithmic NVLink bytes 2 (W-1)/W x message."""
    from iaas_sglang_b200 import tp as tpmod

    dev = cx.dev
    x = torch.randn(rows, hidden, device=dev, dtype=torch.bfloat16)
    res = torch.zeros_like(x)
    w = torch.ones(hidden, device=dev, dtype=torch.bfloat16)
    out = {}
    for name, fn in (("allreduce_us", lambda: tpmod.tensor_model_parallel_all_reduce(x)),
                     ("fused_allreduce_add_rmsnorm_quant_us",
                      lambda: tpmod.tensor_model_parallel_fused_allreduce_rmsnorm(x, res, w, 1e-5, quant_fp8=True, want_half=False))):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                for _ in range(20):
                    fn()
        torch.cuda.synchronize()
        g.replay()
        cx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        cx.barrier()
        out[name] = round(cx.max_over_ranks([e0.elapsed_time(e1) / 20 * 1e3])[0], 2)
        del g
    nbytes = rows * hidden * 2
    out["message_bytes"] = nbytes
    car = tpmod.custom_all_reduce()
    mode = int(car.lib.b200_ar_mode2(cx.world, nbytes, int(car.multicast))) if car is not None else 0
    # received per rank: one-shot push = every peer's whole message; two-shot / NCCL ring = 2 (W-1)/W x message
    out["nvlink_bytes_per_rank"] = int((cx.world - 1) * nbytes if mode == 5 else 2 * (cx.world - 1) / cx.world * nbytes)
    out["kernel"] = {5: "push protocol, one-shot", 6: "push protocol, two-shot"}.get(mode, f"mode {mode}")
    out["path"] = (f"own kernels (csrc/allreduce.cu, {car.setup}, multicast {'yes' if car.multicast else 'no'})"
                   if car is not None and car.should_custom_ar(x) else "ncclAllReduce")
    return out


def tp_check(cx: Ctx, wl: dict):
    """N > 1 gate, run before anything is timed, on two layers of the headline model with every rank's shard cut from
    ONE full set of weights:
      1. emulation: rank 0 recomputes all N shards on its own GPU and combines their bf16 partials the way the all-reduce
         kernels define it (fp32 sum in rank order, one rounding) -- the distributed output (the very code path that is
         timed: own NVLink kernels, fused all-reduce + add + RMSNorm + FP8 quant) must equal it BIT FOR BIT;
      2. every rank holds bit-identical outputs; no all-reduce gave up on a peer (b200_ar_error);
      3. against the UNSHARDED stack: the hidden state after the first reduction (o_proj of layer 0) within 1e-2 -- TP
         only re-orders that sum and rounds the partials to bf16.  The deviation of the final output is reported too; it
         is larger because every later FP8 activation quantisation amplifies a last-bit difference (a code flip is a 6 %
         step), with NCCL + separate kernels just as with ours -- (1) is what pins the kernels."""
    import torch.distributed as dist

    from iaas_sglang_b200 import ops, tp as tpmod

    B, ctx, W = 8 * cx.world, 384, cx.world
    sh = DecodeStack(wl, 2, B, W, cx.rank, cx.dev, collectives=True, shared_weights=True, ctx=ctx)
    sh.probe = {}
    sh.upload()
    sh.backend.init_forward_metadata(sh.fb)
    out = sh.forward().clone()
    mid = sh.probe["residual_after_first_reduce"]
    torch.cuda.synchronize()
    gathered = [torch.empty_like(out) for _ in range(W)]
    dist.all_gather(gathered, out)
    identical = all(torch.equal(gathered[0], g) for g in gathered[1:])
    del sh
    torch.cuda.empty_cache()
    res = torch.zeros(4, device=cx.dev, dtype=torch.float64)   # emulation equal, rel mid, rel final, spare
    if cx.rank == 0:
        # ---- 1. all N shards on this GPU, partials combined in rank order ----
        shards = [DecodeStack(wl, 2, B, W, r, cx.dev, collectives=False, shared_weights=True, ctx=ctx) for r in range(W)]
        for st in shards:
            st.upload()
            st.backend.init_forward_metadata(st.fb)
        residual = torch.zeros_like(shards[0].hidden_in)

        def reduce_norm(parts, w, quant):
            acc = parts[0].float()
            for p_ in parts[1:]:
                acc = acc + p_.float()
            x = acc.to(parts[0].dtype)
            return ops.rmsnorm_quant_fp8(x, w, 1e-5, residual=residual) if quant else ops.rmsnorm(x, w, 1e-5, residual=residual)

        s0 = shards[0]
        x = ops.rmsnorm_quant_fp8(s0.hidden_in, s0.norm1[0], 1e-5, residual=residual)
        for l in range(2):
            if l > 0:
                x = reduce_norm(parts, s0.norm1[l], True)   # noqa: F821
            parts = [st.attn_block(l, x, True) for st in shards]
            x = reduce_norm(parts, s0.norm2[l], True)
            if l == 0:
                emu_mid = residual.clone()
            parts = [st.mlp_block(l, x, True) for st in shards]
        emu = reduce_norm(parts, s0.norm_f, False)
        res[0] = 1.0 if torch.equal(emu, out) else 0.0
        res[3] = ((emu_mid.float() - mid.float()).norm() / emu_mid.float().norm()).item()
        del shards, parts, x, emu
        torch.cuda.empty_cache()
        # ---- 3. the unsharded stack ----
        ref = DecodeStack(wl, 2, B, 1, 0, cx.dev, collectives=False, shared_weights=True, ctx=ctx)
        ref.probe = {}
        ref.upload()
        ref.backend.init_forward_metadata(ref.fb)
        r = ref.forward().float()
        rm = ref.probe["residual_after_first_reduce"].float()
        res[1] = ((mid.float() - rm).norm() / rm.norm()).item()
        res[2] = ((out.float() - r).norm() / r.norm()).item()
        del ref
    dist.broadcast(res, 0)
    car = tpmod.custom_all_reduce()
    err = car.error() if car is not None else 0
    torch.cuda.empty_cache()
    # (NVSwitch multicast reductions add the W copies in the switch's own order: one rounding of the half type around the
    # rank-order sum, tools/ar_check.py -- then the emulation is held to 4e-3 after the first reduction instead of bits)
    nvls = car is not None and car.multicast and W > 2
    emu_ok = res[0].item() == 1.0 or (nvls and 0 <= res[3].item() < 4e-3)
    ok = bool(emu_ok and identical and err == 0 and 0 <= res[1].item() < 1e-2)
    return {"ok": ok, "bit_identical_to_rank_order_emulation": bool(res[0].item() == 1.0),
            "rel_l2_vs_emulation_after_first_reduce": round(res[3].item(), 8), "bit_identical_across_ranks": bool(identical), "ar_error": int(err),
            "rel_l2_vs_unsharded_after_first_reduce": round(res[1].item(), 6), "tolerance": 1e-2,
            "rel_l2_vs_unsharded_final_output": round(res[2].item(), 6), "layers": 2, "batch": B, "ctx": ctx,
            "allreduce": f"own kernels ({car.setup}, multicast {'yes' if car.multicast else 'no'}), fused with add + RMSNorm + FP8 quant"
            if car is not None else "ncclAllReduce"}


def gemm_roofline(dev):
    """'FP8 GEMM TFLOPS vs peak' (BASELINE.json metric): fp8_blockwise_scaled_mm at large M on DeepSeek-V3 weight shapes
    (sgl-kernel/benchmark/bench_fp8_blockwise_gemm.py:21-37) against the dense FP8 peak measured in THIS run the
    MEASURED_PEAKS way (torch._scaled_mm e4m3 8192^3 through cuBLASLt, best of 10).  The kernel-to-beat stand-in
    (vLLM's CUTLASS block-scaled GEMM; the reference's fp8_blockwise_gemm_kernel.cu is the same CUTLASS collective) is
    timed beside ours where it imports."""
    from iaas_sglang_b200 import ops

    def best_ms(fn, n=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(n):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    g = torch.Generator(device=dev).manual_seed(7)
    out = {"peak_source": "torch._scaled_mm e4m3 x e4m3 -> bf16, 8192^3, per-tensor scales (cuBLASLt), best of 10, this run",
           "nominal_dense_fp8_tflops": 4500.0, "kernel": "gemm_big_kernel<fp8 block-scaled>", "cases": []}
    a8 = torch.randn(8192, 8192, device=dev, generator=g).to(torch.float8_e4m3fn)
    w8 = torch.randn(8192, 8192, device=dev, generator=g).to(torch.float8_e4m3fn)
    one = torch.ones((), device=dev)
    try:
        ms = best_ms(lambda: torch._scaled_mm(a8, w8.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16))
        peak = 2.0 * 8192 ** 3 / ms / 1e9
    except Exception as e:  # noqa: BLE001
        peak, out["peak_source"] = 4500.0, f"nominal (torch._scaled_mm failed: {str(e)[:80]})"
    out["peak"] = round(peak, 1)
    del a8, w8
    try:
        from vllm import _custom_ops as vops
    except Exception:  # noqa: BLE001
        vops = None
    shapes = [(8192, 8192, 8192), (4096, 7168, 16384), (4096, 7168, 18432), (4096, 24576, 1536), (4096, 36864, 7168),
              (1024, 7168, 2048), (1024, 4096, 7168)]
    for M, N, K in shapes:
        a = torch.randn(M, K, device=dev, generator=g).to(torch.float8_e4m3fn)
        w = torch.randn(N, K, device=dev, generator=g).to(torch.float8_e4m3fn)
        sa = (torch.rand(K // 128, M, device=dev, generator=g) * 1e-3 + 1e-4).t()
        sb = (torch.rand(N // 128, K // 128, device=dev, generator=g) * 1e-3 + 1e-4).t()
        flops = 2.0 * M * N * K
        ms = best_ms(lambda: ops.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16))
        c = {"M": M, "N": N, "K": K, "tflops": round(flops / ms / 1e9, 1), "frac": round(flops / ms / 1e9 / peak, 4), "us": round(ms * 1e3, 1)}
        if vops is not None:
            try:
                vms = best_ms(lambda: vops.cutlass_scaled_mm(a, w.t(), sa, sb, torch.bfloat16))
                c["vllm_cutlass_blockwise_tflops"] = round(flops / vms / 1e9, 1)
            except Exception:  # noqa: BLE001
                vops = None
        out["cases"].append(c)
        del a, w, sa, sb
    torch.cuda.empty_cache()
    head = out["cases"][0]
    out.update({"M": head["M"], "N": head["N"], "K": head["K"], "tflops": head["tflops"], "frac": head["frac"], "unit": "TFLOP/s",
                "bound": "tensor"})
    return out


def cpu_ops_table(budget_s: float = 20.0):
    """the reference's CPU kernels per op on this box's host cores -- decode_attention_cpu on cfg 1
    exactly (16 x 1024, 32/8/64, page_size 1, shuffled table, 8 splits), fp8_scaled_mm_cpu (prepacked, block 128x128) and
    int8_scaled_mm_with_quant at M in {16, 64}.  Best of 5 after a warm-up."""
    from oracle import ref_ops

    if not ref_ops.available():
        return {"unavailable": "oracle/_ref (the reference's CPU kernels) is not built on this box"}
    ops_ = ref_ops.ops()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(1234)
    bf = torch.bfloat16
    t_start = time.perf_counter()

    def best(fn, n=5):
        fn()
        b = 1e9
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            b = min(b, time.perf_counter() - t0)
        return b * 1e3

    rows = []
    B, ctx, hq, hkv, d = 16, 1024, 32, 8, 64
    T = B * ctx
    kb = torch.randn(T + 1, hkv, d, generator=g).to(bf)
    vb = torch.randn(T + 1, hkv, d, generator=g).to(bf)
    r2t = (torch.randperm(T, generator=g) + 1).view(B, ctx).to(torch.int32)
    q = torch.randn(B, hq, d, generator=g).to(bf)
    kn, vn = torch.randn(B, hkv, d, generator=g).to(bf), torch.randn(B, hkv, d, generator=g).to(bf)
    o = torch.zeros(B, hq, d, dtype=bf)
    logits = torch.zeros(B, hq, 8, d + 1)
    sl, rpi = torch.full((B,), ctx, dtype=torch.int64), torch.arange(B, dtype=torch.int64)
    loc = r2t[:, ctx - 1].to(torch.int64)
    ms = best(lambda: ref_ops.decode_attention_cpu(q, kb, vb, o, kn, vn, loc, logits, r2t, rpi, sl, d ** -0.5, 0.0))
    rows.append({"op": "decode_attention_cpu", "shape": "cfg1: B16 x ctx1024, Hq/Hkv 32/8, D 64, bf16, page_size 1 shuffled",
                 "ms": round(ms, 3), "gbs": round(B * ctx * hkv * d * 2 * 2 / ms / 1e6, 1)})
    for n, k in [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]:
        if time.perf_counter() - t_start > budget_s:
            break
        w8 = torch.randint(-100, 100, (n, k), generator=g, dtype=torch.int8).to(torch.float32).to(torch.float8_e4m3fn)
        ws = torch.rand(n // 128, k // 128, generator=g) * 2e-4 + 1e-4
        wp = ops_.convert_weight_packed(w8)
        wi = torch.randint(-127, 128, (n, k), generator=g, dtype=torch.int8)
        wip = ops_.convert_weight_packed(wi)
        wsc = torch.rand(n, generator=g) * 1e-3
        for M in (16, 64):
            x = (torch.randn(M, k, generator=g) * 0.5).to(bf)
            ms = best(lambda: ops_.fp8_scaled_mm_cpu(x, wp, ws, [128, 128], None, bf, True))
            rows.append({"op": "fp8_scaled_mm_cpu", "shape": f"M{M} N{n} K{k}", "ms": round(ms, 3), "gflops": round(2.0 * M * n * k / ms / 1e6, 1)})
            ms = best(lambda: ops_.int8_scaled_mm_with_quant(x, wip, wsc, None, bf, True))
            rows.append({"op": "int8_scaled_mm_with_quant", "shape": f"M{M} N{n} K{k}", "ms": round(ms, 3), "gops": round(2.0 * M * n * k / ms / 1e6, 1)})
    return {"cores": cores, "kind": "reference", "rows": rows}


# ------------------------------------------------------------------------------------------------
# main
# ------------------------------------------------------------------------------------------------
def extras_for(args, world: int):
    """The other north_star configs measured next to the headline at this N:
    cfg 4 (Llama-70B FP8, shared prefix) at every N, cfg 3 (Qwen-32B AWQ) at N in {1, 2, 4}, cfg 5 (DeepSeek MLA) at 1 and 8."""
    if args.extras == "none" or args.workload != DEFAULT_WORKLOAD or args.tp:
        return []
    if args.extras != "auto":
        return [e for e in args.extras.split(",") if e in WORKLOADS]
    ex = ["cfg4"]
    if world in (1, 2, 4):
        ex.append("cfg3")
    if world in (1, 8):
        ex.append("cfg5")
    return ex


def dtype_string(wl):
    return ("fp8 e4m3 W8A8 GEMMs (fp32 accumulate), bf16 attention (fp32 softmax)" if wl["quant"] == "fp8"
            else f"{wl['quant']} GEMMs, bf16 attention (fp32 softmax)")


def run_extra(cx: Ctx, name: str, args, lib, peak, peak_src):
    """One extra.configs record: the same measurement as the headline (graph replay, CUDA events, max over ranks) on a
    layer slice of another north_star config at TP = N."""
    wl = WORKLOADS[name]
    tp = cx.world
    if wl["hq"] % tp or wl["inter"] % tp or (wl["hkv"] % tp and tp % wl["hkv"]):
        return {"name": name, "skipped": f"heads / widths do not divide by tp{tp}"}
    batch = wl["batch"]
    layers = auto_layers(wl, tp, batch, cap=8 if tp >= 4 else 4)
    t0 = time.time()
    stack = make_stack(wl, layers, batch, tp, cx.rank, cx.dev, collectives=cx.world > 1)
    torch.cuda.synchronize()
    m = measure_stack(cx, stack, max(3, min(args.steps, 10)), 3, e2e=False, lib=lib)
    cfg = workload_config(wl, name, layers, batch, tp)
    if getattr(stack, "cascade", False):
        cfg["attention"] = "cascade decode: shared prefix attended once per batch (extend kernel, prefix_only) + per-request suffix, LSE merge"
    rec = {"name": name, "config": cfg, "dtype": dtype_string(wl),
           "value": round(batch / (m["ms"] / layers * wl["layers"] * 1e-3), 1),
           "unit": f"tok/s ({layers}-layer slice timed, extrapolated linearly over the model's {wl['layers']} identical layers)",
           "slice_ms_per_step": round(m["ms"], 4), "ms_per_layer": round(m["ms"] / layers, 4),
           "launches_per_step": m["launches_per_step"],
           "roofline": roofline_of(stack, m["attn_ms"], peak, peak_src),
           "step_roofline": step_roofline_of(stack, m["ms"], peak)}
    if wl["shared_prefix"] and wl.get("arch") != "mla" and not stack.cascade:
        rec["roofline"]["cascade"] = measure_cascade_attention(cx, stack, rec["roofline"]["ms_per_launch"])
    if cx.world > 1:
        rec["allreduce"] = time_allreduce(cx, batch, wl["hidden"])
    rec["wall_s"] = round(time.time() - t0, 1)
    del stack
    torch.cuda.empty_cache()
    return rec


def main():
    args = parse_args()
    claim_stdout()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl, args.layers or wl["layers"])
        return

    import torch.distributed as dist

    from iaas_sglang_b200 import _lib, tp as tpmod

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != max(1, args.gpus):
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback); use --impl reference for the CPU arm"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        tpmod.init_distributed("nccl", dev)
    lib = _lib.load()
    assert lib.b200_sm_arch() == 100, f"sm_{lib.b200_sm_arch()} is not a B200"
    cx = Ctx(world, rank, dev)
    peak, peak_src = measured_peaks()
    t_run = time.time()

    # ---- N > 1 first gate: the sharded stack equals the unsharded one (same weights), identical on every rank ----
    check = None
    if world > 1 and not args.no_tp_check and wl["quant"] == "fp8" and wl.get("arch") != "mla":
        check = tp_check(cx, wl)
        log(f"[bench] tp_check {check}")
        if not check["ok"]:
            if rank == 0:
                emit({"metric": "decode tok/s", "value": None, "n_gpus": world, "tp_check": check,
                      "error": "tensor-parallel output failed the parity gate: nothing was timed"})
            sys.stdout.flush()
            os._exit(3)

    tp = args.tp or world
    batch = args.batch or wl["batch"] * (world if wl.get("scale_batch") else 1)
    n_layers = args.layers or (wl["layers"] if args.workload.startswith("cfg2") else auto_layers(wl, tp, batch, cap=8))
    stack = make_stack(wl, n_layers, batch, tp, rank if world > 1 else 0, dev, collectives=world > 1)
    stack.no_fused = args.no_fused
    torch.cuda.synchronize()

    clocks = ClockSampler(local)
    clocks.start()
    m = measure_stack(cx, stack, args.steps, args.warmup, e2e=not args.no_e2e, lib=lib)
    clk = clocks.stop()
    ms, e2e_ms, attn_ms = m["ms"], m["e2e_ms"], m["attn_ms"]

    if args.kernels and hasattr(stack.method, "apply_quantized") and wl.get("arch") != "mla":
        kernels_breakdown(stack, ms, attn_ms, batch, n_layers, dev)

    traffic = traffic_src = None
    try:   # per-launch DRAM bytes of the same kernel at the same shape, from the committed ncu capture (NOT this run)
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            ent = json.load(f).get(f"{args.workload}/tp{tp}")
        if ent and batch == wl["batch"] * (tp if wl.get("scale_batch") else 1):
            traffic = int(ent["bytes"])
            traffic_src = f"ncu --set full capture committed under profiles/ ({ent.get('source', 'traffic.json')}), not measured in this run"
    except (OSError, ValueError, KeyError):
        pass

    line = None
    if rank == 0:
        value = batch / (ms * 1e-3)
        line = {
            "metric": "decode tok/s", "value": round(value, 1), "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": dtype_string(wl),
            "data": "synthetic (random-init weights, randn KV, shuffled page table, seed 1234)",
            "config": workload_config(wl, args.workload, n_layers, batch, tp),
            "e2e": None if e2e_ms is None else {
                "value": round(batch / (e2e_ms * 1e-3), 1), "unit": "tok/s", "ms_per_step": round(e2e_ms, 4),
                "h2d_bytes_per_step": stack.h2d_bytes(), "d2h_bytes_per_step": stack.d2h_bytes(),
                "path": "pinned host -> static graph buffers -> CUDA-graph replay of B200AttnBackend/linear methods -> pinned host, sync per step"},
            "gpu_launches": m["launches_per_step"] * args.steps,
            "launches_per_step": m["launches_per_step"],
            "clocks": clk,
            "roofline": roofline_of(stack, attn_ms, peak, peak_src, traffic, traffic_src),
            "step_roofline": step_roofline_of(stack, ms, peak),
        }
        if check is not None:
            line["tp_check"] = check
    if world > 1:
        ar = time_allreduce(cx, batch, wl["hidden"])
        if rank == 0:
            line["allreduce"] = ar
            car = tpmod.custom_all_reduce()
            line["config"]["allreduce"] = ("fused all-reduce + residual-add + RMSNorm + FP8 quant, own NVLink kernels (csrc/allreduce.cu), error word %d"
                                           % car.error() if car is not None else "ncclAllReduce")
    del stack
    torch.cuda.empty_cache()

    # ---- the other north_star configs at this N ----
    extras = []
    for name in extras_for(args, world):
        if time.time() - t_run > 240:
            extras.append({"name": name, "skipped": "time budget of the default run"})
            continue
        try:
            rec = run_extra(cx, name, args, lib, peak, peak_src)
        except torch.cuda.OutOfMemoryError as e:  # noqa: PERF203
            rec = {"name": name, "skipped": f"out of memory: {str(e)[:80]}"}
            torch.cuda.empty_cache()
        log(f"[bench] extra {name}: {json.dumps(rec)[:600]}")
        extras.append(rec)

    if rank == 0:
        line["extra"] = {"configs": extras}
        # ---- FP8 GEMM TFLOPS vs the dense FP8 peak (one third of BASELINE.json's metric) ----
        if not args.no_gemm_roofline and world == 1:
            try:
                line["roofline_gemm"] = gemm_roofline(dev)
            except Exception as e:  # noqa: BLE001
                line["roofline_gemm"] = {"error": str(e)[:200]}
        if not args.no_cpu_baseline and world == 1 and wl["quant"] == "fp8" and wl.get("arch") != "mla":
            try:
                cores = os.cpu_count() or 1
                sample = CpuLayerSample(wl, batch, cores)
                val, cms, done = time_cpu_layers(sample, 1, 1000, n_layers, budget_s=12.0)
                line["cpu_baseline"] = {"value": round(val, 3), "unit": "tok/s", "cores": sample.threads, "kind": sample.kind,
                                        "sample": sample.describe(n_layers) + f"; {done} timed layer passes, {cms:.1f} ms each"}
                del sample
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "port", "sample": f"failed: {e}"}
            try:
                line["extra"]["cpu_ops"] = cpu_ops_table()
            except Exception as e:  # noqa: BLE001
                line["extra"]["cpu_ops"] = {"error": str(e)[:200]}
        line["wall_s"] = round(time.time() - t_run, 1)
        emit(line)
    if world > 1:
        # NCCL teardown with captured collectives still alive can block for minutes (seen on the 2-GPU box: the JSON
        # line was out, destroy_process_group never returned).  The measurement is complete: leave through a last
        # barrier with a watchdog, and skip the communicator teardown.
        sys.stdout.flush()
        watchdog = threading.Timer(20.0, lambda: os._exit(0))
        watchdog.daemon = True
        watchdog.start()
        try:
            torch.cuda.synchronize()
            dist.barrier()
        finally:
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(0)


def kernels_breakdown(stack, ms, attn_ms, batch, n_layers, dev):
    """developer aid (stderr): the four linears of every layer alone, as one graph"""
    from iaas_sglang_b200 import ops

    def xq(kdim):
        return ops.per_token_group_quant_fp8(torch.randn(batch, kdim, device=dev, dtype=torch.bfloat16) / 10, 128,
                                             column_major_scales=True)
    x_h, x_a, x_i = xq(stack.hidden), xq(stack.hq * stack.d), xq(stack.inter)

    def gemms():
        for l in range(n_layers):
            stack.method.apply_quantized(stack.w_qkv[l], x_h[0], x_h[1], stack.dtype)
            stack.method.apply_quantized(stack.w_o[l], x_a[0], x_a[1], stack.dtype)
            stack.method.apply_quantized(stack.w_gate_up[l], x_h[0], x_h[1], stack.dtype)
            stack.method.apply_quantized(stack.w_down[l], x_i[0], x_i[1], stack.dtype)
    sg = torch.cuda.Stream()
    sg.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(sg):
        gemms()
    torch.cuda.current_stream().wait_stream(sg)
    gg = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gg, stream=sg):
        gemms()
    for _ in range(3):
        gg.replay()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(10):
        gg.replay()
    g1.record()
    torch.cuda.synchronize()
    gemm_us = g0.elapsed_time(g1) / 10 / n_layers * 1e3
    log(json.dumps({"per_layer_us": {"step": round(ms / n_layers * 1e3, 1), "attention+merge": round(attn_ms * 1e3, 1),
                                     "four_linears_alone": round(gemm_us, 1),
                                     "rest (norm+quant x2, rope, silu*mul+quant, gaps)": round(ms / n_layers * 1e3 - attn_ms * 1e3 - gemm_us, 1)},
                    "weight_bytes_per_layer": stack.weight_bytes_per_layer()}))


if __name__ == "__main__":
    main()
